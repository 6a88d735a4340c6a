// Hessian-free building block for the scaled logistic configs (BASELINE.json config 3: 512 workers
// x d = 8192, local Newton solves): one pass over the shard gives  v = X' w(X x)  where
//   GRAD mode:  z = X x,  w = -y ./ (1 + exp(y .* z)),  also stores D = sigma(1-sigma) and the loss
//               sum log(1 + exp(-y .* z))        -> v is the data part of the gradient
//   HV   mode:  u = X p,  w = D .* u              -> v = X' D X p (Gauss-Newton/Hessian product)
// With s < d an explicit d x d Hessian (2 s d^2 flop + d^3 for the factorisation, per Newton step
// and worker) is far more work than a handful of these products (4 s d flop, one read of X each),
// so the exact local solve of group_ADMM_logistic.m:48-55 is done by Newton-CG.
//
// Kernel: grid (G, N).  CTA (g, n) owns rows g, g+G, ... of worker n's shard (fp32, row-major).
// Rows are streamed in blocks of RB rows with cp.async.bulk into a double-buffered shared tile, so
// each row is read from HBM exactly once per product: first the RB dot products with x (x lives in
// shared memory, fp64), then the rank-RB update of the column sums, which stay in registers
// (each thread owns d/256 columns) for the CTA's whole lifetime.  The G partial column sums per
// worker are added in a fixed order by a second tiny kernel (deterministic).
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

constexpr int LS_THREADS = 256;
constexpr int LS_MAXC = 32;         // columns per thread -> d <= 8192

// LS_RB: rows per streamed block (8/4/2, the largest whose two buffers fit next to x).
// FAST (hv == 2 only): products in fp32 (x and the row weights rounded to fp32), fp64 accumulation
// across threads / row blocks -- one F2F.F64 per thread-partial instead of two per element, which
// is what bounds the exact path.  ~1e-6 relative, for the inner CG only; gradients stay exact.
template <int LS_RB, bool FAST>
__global__ void __launch_bounds__(LS_THREADS, 1)
logistic_xtwx_kernel(const float* __restrict__ X, const double* __restrict__ y,
                     const double* __restrict__ xin, long long x_stride, double* __restrict__ Dw,
                     double* __restrict__ part, double* __restrict__ losspart, int s, int d, int hv,
                     int shard_stride) {
  extern __shared__ __align__(128) unsigned char dsm[];
  const int n = blockIdx.y, g = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  float* tile = reinterpret_cast<float*>(dsm);                         // [2][RB][d]
  double* xs = reinterpret_cast<double*>(dsm + (size_t)2 * LS_RB * d * sizeof(float));   // [d]
  double* wr = xs + d;                                                 // [RB] row weights
  double* dots = wr + LS_RB;                                           // [RB][8] per-warp partials
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ double red[32];

  const size_t shard = (size_t)n * shard_stride;           // e.g. stride 2 = one colour of a chain
  const float* Xn = X + shard * s * d;
  const double* yn = y + shard * s;
  double* Dn = Dw + shard * s;
  float* xs32 = reinterpret_cast<float*>(xs);
  float* wr32 = reinterpret_cast<float*>(wr);
  for (int j = tid; j < d; j += LS_THREADS) {
    const double xv = xin[(size_t)n * x_stride + j];
    if constexpr (FAST) xs32[j] = (float)xv; else xs[j] = xv;
  }
  if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_fence_init(); }
  __syncthreads();

  // thread tid owns columns 4*tid + 1024*c + {0,1,2,3}, c = 0..7 (128-bit shared-memory loads)
  double acc[LS_MAXC];
#pragma unroll
  for (int c = 0; c < LS_MAXC; ++c) acc[c] = 0.0;
  // fp64 modes are bound by the fp32->fp64 conversions (quarter-rate pipe).  With two rows per
  // block the converted block fits in registers between the two phases (a thread touches the same
  // columns in both), which halves the conversions and the shared-memory reads.
  constexpr bool KEEP = !FAST && LS_RB == 2;
  double keep[KEEP ? LS_RB : 1][KEEP ? LS_MAXC : 1];
  double loss = 0.0;
  const int nblk = (s + LS_RB - 1) / LS_RB;                 // row blocks of the shard
  const int mine = (nblk - g + G - 1) / G;                  // blocks g, g+G, ... handled here
  const uint32_t rowbytes = (uint32_t)d * sizeof(float);

  auto issue = [&](int k) {                                 // k-th block of this CTA -> buffer k&1
    const int r0 = (g + k * G) * LS_RB;
    const int rows = min(LS_RB, s - r0);
    mbar_arrive_expect_tx(&bar[k & 1], rows * rowbytes);
    bulk_g2s(tile + (size_t)(k & 1) * LS_RB * d, Xn + (size_t)r0 * d, rows * rowbytes, &bar[k & 1]);
  };
  if (tid == 0 && mine > 0) issue(0);
  for (int k = 0; k < mine; ++k) {
    if (tid == 0 && k + 1 < mine) issue(k + 1);
    const int r0 = (g + k * G) * LS_RB;
    const int rows = min(LS_RB, s - r0);
    // the row's label / curvature weight is needed only after the dot products: fetch it now
    double rowval = 0.0;
    if (tid < rows) rowval = hv ? Dn[r0 + tid] : yn[r0 + tid];
    mbar_wait(&bar[k & 1], (k >> 1) & 1);
    const float* t = tile + (size_t)(k & 1) * LS_RB * d;
    // ---- dot products of the block's rows with x
    double dp[LS_RB];
    if constexpr (FAST) {
      float dpf[LS_RB];
#pragma unroll
      for (int r = 0; r < LS_RB; ++r) dpf[r] = 0.f;
#pragma unroll
      for (int c = 0; c < LS_MAXC / 4; ++c) {
        const int j = 4 * tid + 4 * LS_THREADS * c;
        if (j < d) {
          const float4 xv = *reinterpret_cast<const float4*>(xs32 + j);
#pragma unroll
          for (int r = 0; r < LS_RB; ++r) {
            const float4 tv = *reinterpret_cast<const float4*>(t + (size_t)r * d + j);
            dpf[r] = fmaf(tv.x, xv.x, dpf[r]);
            dpf[r] = fmaf(tv.y, xv.y, dpf[r]);
            dpf[r] = fmaf(tv.z, xv.z, dpf[r]);
            dpf[r] = fmaf(tv.w, xv.w, dpf[r]);
          }
        }
      }
#pragma unroll
      for (int r = 0; r < LS_RB; ++r) dp[r] = (double)dpf[r];
    } else {
#pragma unroll
      for (int r = 0; r < LS_RB; ++r) dp[r] = 0.0;
#pragma unroll
      for (int c = 0; c < LS_MAXC / 4; ++c) {
        const int j = 4 * tid + 4 * LS_THREADS * c;
        if (j < d) {
          const double2 xa = *reinterpret_cast<const double2*>(xs + j);
          const double2 xb = *reinterpret_cast<const double2*>(xs + j + 2);
#pragma unroll
          for (int r = 0; r < LS_RB; ++r) {
            const float4 tv = *reinterpret_cast<const float4*>(t + (size_t)r * d + j);
            const double t0 = (double)tv.x, t1 = (double)tv.y, t2 = (double)tv.z, t3 = (double)tv.w;
            if constexpr (KEEP) {       // the column phase needs the same elements: convert once
              keep[r][4 * c + 0] = t0; keep[r][4 * c + 1] = t1;
              keep[r][4 * c + 2] = t2; keep[r][4 * c + 3] = t3;
            }
            dp[r] = fma(t0, xa.x, dp[r]);
            dp[r] = fma(t1, xa.y, dp[r]);
            dp[r] = fma(t2, xb.x, dp[r]);
            dp[r] = fma(t3, xb.y, dp[r]);
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < LS_RB; ++r) {
      dp[r] = warp_sum(dp[r]);
      if (lane == 0) dots[r * 8 + warp] = dp[r];
    }
    __syncthreads();
    if (tid < LS_RB) {
      double z = 0.0;
#pragma unroll
      for (int w8 = 0; w8 < 8; ++w8) z += dots[tid * 8 + w8];
      double wv = 0.0;
      if (tid < rows) {
        if (hv) {
          wv = rowval * z;
        } else {
          const double yz = rowval * z;
          const double e = exp(-fabs(yz));                     // overflow-free forms
          const double sg = yz >= 0.0 ? e / (1.0 + e) : 1.0 / (1.0 + e);   // 1/(1+exp(yz))
          wv = -rowval * sg;
          Dn[r0 + tid] = sg * (1.0 - sg);
          loss += fmax(-yz, 0.0) + log1p(e);
        }
      }
      if constexpr (FAST) wr32[tid] = (float)wv; else wr[tid] = wv;
    }
    __syncthreads();
    // ---- column sums: acc += sum_r t[r][j] * w_r (rows beyond the shard are stale shared memory)
#pragma unroll
    for (int c = 0; c < LS_MAXC / 4; ++c) {
      const int j = 4 * tid + 4 * LS_THREADS * c;
      if (j < d) {
        if constexpr (FAST) {
          float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
          for (int r = 0; r < LS_RB; ++r)
            if (r < rows) {
              const float4 tv = *reinterpret_cast<const float4*>(t + (size_t)r * d + j);
              const float w = wr32[r];
              a0 = fmaf(tv.x, w, a0); a1 = fmaf(tv.y, w, a1);
              a2 = fmaf(tv.z, w, a2); a3 = fmaf(tv.w, w, a3);
            }
          acc[4 * c + 0] += (double)a0; acc[4 * c + 1] += (double)a1;
          acc[4 * c + 2] += (double)a2; acc[4 * c + 3] += (double)a3;
        } else if constexpr (KEEP) {
#pragma unroll
          for (int r = 0; r < LS_RB; ++r)
            if (r < rows) {
              const double w = wr[r];
              acc[4 * c + 0] = fma(keep[r][4 * c + 0], w, acc[4 * c + 0]);
              acc[4 * c + 1] = fma(keep[r][4 * c + 1], w, acc[4 * c + 1]);
              acc[4 * c + 2] = fma(keep[r][4 * c + 2], w, acc[4 * c + 2]);
              acc[4 * c + 3] = fma(keep[r][4 * c + 3], w, acc[4 * c + 3]);
            }
        } else {
#pragma unroll
          for (int r = 0; r < LS_RB; ++r)
            if (r < rows) {
              const float4 tv = *reinterpret_cast<const float4*>(t + (size_t)r * d + j);
              const double w = wr[r];
              acc[4 * c + 0] = fma((double)tv.x, w, acc[4 * c + 0]);
              acc[4 * c + 1] = fma((double)tv.y, w, acc[4 * c + 1]);
              acc[4 * c + 2] = fma((double)tv.z, w, acc[4 * c + 2]);
              acc[4 * c + 3] = fma((double)tv.w, w, acc[4 * c + 3]);
            }
        }
      }
    }
    __syncthreads();                                        // buffer k&1 and wr are free again
  }
  double* pn = part + ((size_t)n * G + g) * d;
#pragma unroll
  for (int c = 0; c < LS_MAXC / 4; ++c) {
    const int j = 4 * tid + 4 * LS_THREADS * c;
    if (j < d) {
      *reinterpret_cast<double2*>(pn + j) = make_double2(acc[4 * c + 0], acc[4 * c + 1]);
      *reinterpret_cast<double2*>(pn + j + 2) = make_double2(acc[4 * c + 2], acc[4 * c + 3]);
    }
  }
  if (!hv) {
    loss = block_sum(tid < LS_RB ? loss : 0.0, red);
    if (tid == 0) losspart[(size_t)n * G + g] = loss;
  }
}

// v[n][j] = sum_g part[n][g][j] (fixed order), loss[n] = sum_g losspart[n][g]
__global__ void __launch_bounds__(256)
logistic_reduce_kernel(const double* __restrict__ part, const double* __restrict__ losspart,
                       double* __restrict__ v, double* __restrict__ loss, int G, int d, int ldv, int hv) {
  const int n = blockIdx.y;
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < d) {
    double a = 0.0;
    for (int g = 0; g < G; ++g) a += part[((size_t)n * G + g) * d + j];
    v[(size_t)n * ldv + j] = a;
  }
  if (!hv && blockIdx.x == 0 && threadIdx.x == 0) {
    double l = 0.0;
    for (int g = 0; g < G; ++g) l += losspart[(size_t)n * G + g];
    loss[n] = l;
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

size_t gadmm_logistic_xtwx_scratch_elems(int N, int d, int G) { return (size_t)N * G * (d + 1); }

/* v[n] = X_n' w(X_n x_n) for N shards X: [N][s][d] fp32, y: [N][s] (+-1, fp64), x: fp64 with row
 * stride x_stride (0 = one shared vector), Dw: [N][s] (written in GRAD mode, read in HV mode),
 * v: [N][ldv], loss: [N] (GRAD mode).  scratch: gadmm_logistic_xtwx_scratch_elems doubles.
 * Shard n of the call is shard n*shard_stride of X / y / Dw (x, v, loss are dense over n).
 * d must be a multiple of 4 (16-byte rows) and <= 8192. */
int gadmm_logistic_xtwx(const float* X, const double* y, const double* x, long long x_stride,
                        double* Dw, double* v, int ldv, double* loss, int N, int s, int d, int G,
                        int hv, int shard_stride, double* scratch, void* stream) {
  GADMM_REQUIRE(N > 0 && s > 0 && d > 0 && G > 0 && hv >= 0 && hv <= 2, "logistic_xtwx: bad shape");
  GADMM_REQUIRE(d % 4 == 0 && d <= LS_MAXC * LS_THREADS,
                "logistic_xtwx: d=%d must be a multiple of 4 and <= %d", d, LS_MAXC * LS_THREADS);
  const int rb = (size_t)d * 72 <= 200 * 1024 ? 8 : ((size_t)d * 40 <= 200 * 1024 ? 4 : 2);
  const size_t smem = (size_t)2 * rb * d * sizeof(float) + (size_t)(d + rb + rb * 8) * sizeof(double);
  GADMM_REQUIRE(smem <= 220 * 1024, "logistic_xtwx: d=%d needs %zu B of shared memory", d, smem);
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (!configured) {
    const int cap = 220 * 1024;
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_xtwx_kernel<8, false>, attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_xtwx_kernel<4, false>, attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_xtwx_kernel<2, false>, attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_xtwx_kernel<8, true>, attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_xtwx_kernel<4, true>, attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_xtwx_kernel<2, true>, attr, cap));
    configured = 1;
  }
  cudaStream_t st = (cudaStream_t)stream;
  double* part = scratch;
  double* losspart = scratch + (size_t)N * G * d;
  const dim3 grid(G, N);
#define GADMM_XTWX_LAUNCH(RB, FAST)                                                              \
  logistic_xtwx_kernel<RB, FAST><<<grid, LS_THREADS, smem, st>>>(X, y, x, x_stride, Dw, part,    \
                                                                losspart, s, d, hv, shard_stride)
  const bool fast = hv == 2;
  if (rb == 8) { if (fast) GADMM_XTWX_LAUNCH(8, true); else GADMM_XTWX_LAUNCH(8, false); }
  else if (rb == 4) { if (fast) GADMM_XTWX_LAUNCH(4, true); else GADMM_XTWX_LAUNCH(4, false); }
  else { if (fast) GADMM_XTWX_LAUNCH(2, true); else GADMM_XTWX_LAUNCH(2, false); }
#undef GADMM_XTWX_LAUNCH
  logistic_reduce_kernel<<<dim3((d + 255) / 256, N), 256, 0, st>>>(part, losspart, v, loss, G, d, ldv,
                                                                   hv);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
