// Device-side topology + random data (K10, K13, K14 of SURVEY.md section 2.4).
//
//   philox_uniform / philox_normal   counter-based Philox4x32-10 fills (seeded, reproducible;
//                                    the reference uses unseeded rand/randn)
//   find_path_kernel                 findPath.m:16-30 / findPath2.m:33-59: greedy nearest-neighbour
//                                    chain from node 1 (one CTA, block-wide argmin per hop, first
//                                    minimum wins, strict '<' against the 1e12 sentinel), hop costs
//                                    d^2 or the link-energy model P = d^2*eta*B*2^(R/B), the star
//                                    model P_central and the centre node (findPath2.m:16-29)
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

struct Philox {
  uint32_t key[2];
  __device__ __forceinline__ static void round(uint32_t c[4], const uint32_t k[2]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
    const uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k[0], n1 = lo1, n2 = hi0 ^ c[3] ^ k[1], n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
  }
  __device__ __forceinline__ static void gen(uint64_t seed, uint64_t ctr, uint32_t stream,
                                             uint32_t out[4]) {
    uint32_t k[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t c[4] = {(uint32_t)ctr, (uint32_t)(ctr >> 32), stream, 0x9E3779B9u};
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      round(c, k);
      k[0] += 0x9E3779B9u;
      k[1] += 0xBB67AE85u;
    }
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
  }
};

// (dx^2 + dy^2) without FMA contraction, so that ties and costs match the host's float64 exactly
__device__ __forceinline__ double sq2(double dx, double dy) {
  return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}

__device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
  const uint64_t v = (((uint64_t)a << 32) | b) >> 11;
  return (double)v * (1.0 / 9007199254740992.0);           // [0,1)
}

__global__ void philox_uniform_kernel(double* out, size_t n, uint64_t seed, uint32_t stream,
                                      double scale) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i * 2 < n;
       i += (size_t)gridDim.x * blockDim.x) {
    uint32_t r[4];
    Philox::gen(seed, i, stream, r);
    out[2 * i] = scale * u53(r[0], r[1]);
    if (2 * i + 1 < n) out[2 * i + 1] = scale * u53(r[2], r[3]);
  }
}

template <typename T>
__global__ void philox_normal_kernel(T* out, size_t n, uint64_t seed, uint32_t stream, double mean,
                                     double sd) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i * 2 < n;
       i += (size_t)gridDim.x * blockDim.x) {
    uint32_t r[4];
    Philox::gen(seed, i, stream, r);
    const double u1 = 1.0 - u53(r[0], r[1]), u2 = u53(r[2], r[3]);     // u1 in (0,1]
    const double rad = sqrt(-2.0 * log(u1));
    double sn, cs;
    sincospi(2.0 * u2, &sn, &cs);
    out[2 * i] = (T)(mean + sd * rad * cs);
    if (2 * i + 1 < n) out[2 * i + 1] = (T)(mean + sd * rad * sn);
  }
}

// ------------------------------------------------------------------------------ findPath
// xy: [N][2] node positions.  model 0: findPath (cost = d^2); model 1: findPath2 (cost = P).
__global__ void __launch_bounds__(1024)
find_path_kernel(const double* __restrict__ xy, int N, int model, double area, int* __restrict__ path,
                 double* __restrict__ cost, double* __restrict__ d_square,
                 double* __restrict__ P_central, int* __restrict__ center) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double* px = reinterpret_cast<double*>(dsm);
  double* py = px + N;
  unsigned char* visited = reinterpret_cast<unsigned char*>(py + N);
  __shared__ double s_val[32];
  __shared__ int s_idx[32];
  __shared__ int s_cur;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  // batched: CTA b builds chain b from xy[b] (d_square / P_central / center only for a single chain)
  xy += (size_t)blockIdx.x * 2 * N;
  path += (size_t)blockIdx.x * N;
  cost += (size_t)blockIdx.x * (N > 1 ? N - 1 : 1);
  for (int i = tid; i < N; i += blockDim.x) {
    px[i] = xy[2 * i];
    py[i] = xy[2 * i + 1];
    visited[i] = 0;
  }
  __syncthreads();
  if (d_square)
    for (int i = tid; i < N * N; i += blockDim.x) {
      const int a = i / N, b = i % N;
      const double dx = px[a] - px[b], dy = py[a] - py[b];
      d_square[i] = (a == b) ? 0.0 : sq2(dx, dy);
    }
  auto block_argmin = [&](double v, int idx) -> int {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ov < v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    if (lane == 0) { s_val[warp] = v; s_idx[warp] = idx; }
    __syncthreads();
    if (warp == 0) {
      v = lane < nw ? s_val[lane] : INFINITY;
      idx = lane < nw ? s_idx[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov < v || (ov == v && oi < idx)) { v = ov; idx = oi; }
      }
      if (lane == 0) { s_idx[0] = idx; s_val[0] = v; }
    }
    __syncthreads();
    const int r = s_idx[0];
    __syncthreads();
    return r;
  };
  if (model == 1 && center) {                                         // findPath2.m:16-29
    double best = INFINITY; int bi = 0x7fffffff;
    for (int i = tid; i < N; i += blockDim.x) {
      const double dx = px[i] - area / 2, dy = py[i] - area / 2;
      const double dd = sqrt(sq2(dx, dy));
      if (dd < best) { best = dd; bi = i; }
    }
    const int c = block_argmin(best, bi);
    if (tid == 0) *center = c;
    if (P_central)
      for (int i = tid; i < N; i += blockDim.x) {
        const double dx = px[i] - px[c], dy = py[i] - py[c];
        P_central[i] = __dmul_rn(__dmul_rn(__dmul_rn(__dmul_rn(0.5, sq2(dx, dy)), 1e-6), 2e6), exp2(2.0 * 10e6 / 2e6));
      }
  }
  if (tid == 0) { s_cur = 0; visited[0] = 1; path[0] = 0; }
  __syncthreads();
  for (int step = 1; step < N; ++step) {
    const int cur = s_cur;
    double best = INFINITY; int bi = 0x7fffffff;
    for (int m = tid; m < N; m += blockDim.x) {
      if (visited[m] || m == cur) continue;
      const double dx = px[cur] - px[m], dy = py[cur] - py[m];
      const double dd = sq2(dx, dy);
      if (dd < 1e12 && dd < best) { best = dd; bi = m; }              // strict '<', 1e12 sentinel
    }
    const int nxt = block_argmin(best, bi);
    if (tid == 0) {
      const double dx = px[cur] - px[nxt], dy = py[cur] - py[nxt];
      const double dd = sq2(dx, dy);
      path[step] = nxt;
      // P(n,m) = d^2*eta*B*2^(R/B), evaluated left to right like the reference (findPath2.m:37)
      cost[step - 1] = model == 1 ? __dmul_rn(__dmul_rn(__dmul_rn(dd, 1e-6), 2e6), exp2(10e6 / 2e6)) : dd;
      visited[nxt] = 1;
      s_cur = nxt;
    }
    __syncthreads();
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

int gadmm_philox_uniform(double* out, size_t n, unsigned long long seed, unsigned int stream_id,
                         double scale, void* stream) {
  if (n == 0) return 0;
  const int blocks = (int)((n / 2 + 255) / 256 < 2048 ? (n / 2 + 255) / 256 + 1 : 2048);
  philox_uniform_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(out, n, seed, stream_id, scale);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_philox_normal(void* out, size_t n, int is_f32, unsigned long long seed,
                        unsigned int stream_id, double mean, double sd, void* stream) {
  if (n == 0) return 0;
  const int blocks = (int)((n / 2 + 255) / 256 < 4096 ? (n / 2 + 255) / 256 + 1 : 4096);
  if (is_f32)
    philox_normal_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>((float*)out, n, seed,
                                                                         stream_id, mean, sd);
  else
    philox_normal_kernel<double><<<blocks, 256, 0, (cudaStream_t)stream>>>((double*)out, n, seed,
                                                                          stream_id, mean, sd);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_find_path_batch(const double* xy, int K, int N, int model, int* path, double* cost,
                          void* stream);

int gadmm_find_path(const double* xy, int N, int model, int* path, double* cost, double* d_square,
                    double* P_central, int* center, void* stream) {
  GADMM_REQUIRE(N >= 1 && N <= 8192, "find_path: N=%d out of range [1, 8192]", N);
  GADMM_REQUIRE(model == 0 || model == 1, "find_path: model must be 0 (findPath) or 1 (findPath2)");
  const int smem = N * (2 * (int)sizeof(double) + 1) + 16;
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (smem > 48 * 1024 && smem > configured) {
    GADMM_CUDA_OK(cudaFuncSetAttribute(find_path_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = smem;
  }
  find_path_kernel<<<1, 1024, smem, (cudaStream_t)stream>>>(xy, N, model, model == 1 ? 250.0 : 50.0,
                                                          path, cost, d_square, P_central, center);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}


/* K chains at once: xy [K][N][2] -> path [K][N] (0-based), cost [K][max(N-1,1)]; one CTA each. */
int gadmm_find_path_batch(const double* xy, int K, int N, int model, int* path, double* cost,
                          void* stream) {
  GADMM_REQUIRE(N >= 1 && N <= 8192 && K >= 1, "find_path_batch: bad N=%d K=%d", N, K);
  GADMM_REQUIRE(model == 0 || model == 1, "find_path_batch: model must be 0 or 1");
  const int smem = N * (2 * (int)sizeof(double) + 1) + 16;
  if (smem > 48 * 1024)
    GADMM_CUDA_OK(cudaFuncSetAttribute(find_path_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  find_path_kernel<<<K, 1024, smem, (cudaStream_t)stream>>>(xy, N, model, model == 1 ? 250.0 : 50.0,
                                                          path, cost, nullptr, nullptr, nullptr);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
