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

#include <cooperative_groups.h>

namespace gadmm {

constexpr int LS_THREADS = 256;
constexpr int LS_MAXC = 32;         // columns per thread -> d <= 8192

// LS_RB: rows per streamed block (8/4/2, the largest whose two buffers fit next to x).
// FAST (hv == 2 only): products in fp32 (x and the row weights rounded to fp32), fp64 accumulation
// across threads / row blocks -- one F2F.F64 per thread-partial instead of two per element, which
// is what bounds the exact path.  ~1e-6 relative, for the inner CG only; gradients stay exact.
// 4 consecutive shard elements as doubles (fp32 shards are widened here, fp64 shards are exact)
__device__ __forceinline__ void ld4(const float* p, double& a, double& b, double& c, double& e) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  a = (double)v.x; b = (double)v.y; c = (double)v.z; e = (double)v.w;
}
__device__ __forceinline__ void ld4(const double* p, double& a, double& b, double& c, double& e) {
  const double2 u = *reinterpret_cast<const double2*>(p);
  const double2 v = *reinterpret_cast<const double2*>(p + 2);
  a = u.x; b = u.y; c = v.x; e = v.y;
}

// TX = float: fp32 shards (half the HBM bytes; data that is exactly representable in fp32, e.g.
// the synthetic configs).  TX = double: fp64 shards, so that user data is never rounded
// (group_ADMM_logistic.m computes in double throughout).
template <typename TX, int LS_RB, bool FAST>
__global__ void __launch_bounds__(LS_THREADS, 1)
logistic_xtwx_kernel(const void* __restrict__ Xv, const double* __restrict__ y,
                     const double* __restrict__ xin, long long x_stride, double* __restrict__ Dw,
                     double* __restrict__ part, double* __restrict__ losspart, int s, int d, int hv,
                     int shard_stride) {
  extern __shared__ __align__(128) unsigned char dsm[];
  const int n = blockIdx.y, g = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  static_assert(!FAST || sizeof(TX) == 4, "fp32-multiply mode needs fp32 shards");
  TX* tile = reinterpret_cast<TX*>(dsm);                               // [2][RB][d]
  double* xs = reinterpret_cast<double*>(dsm + (size_t)2 * LS_RB * d * sizeof(TX));      // [d]
  double* wr = xs + d;                                                 // [RB] row weights
  double* dots = wr + LS_RB;                                           // [RB][8] per-warp partials
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ double red[32];

  const size_t shard = (size_t)n * shard_stride;           // e.g. stride 2 = one colour of a chain
  const TX* Xn = reinterpret_cast<const TX*>(Xv) + shard * s * d;
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
  constexpr bool KEEP = !FAST && LS_RB <= 2;
  double keep[KEEP ? LS_RB : 1][KEEP ? LS_MAXC : 1];
  double loss = 0.0;
  const int nblk = (s + LS_RB - 1) / LS_RB;                 // row blocks of the shard
  const int mine = (nblk - g + G - 1) / G;                  // blocks g, g+G, ... handled here
  const uint32_t rowbytes = (uint32_t)d * sizeof(TX);

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
    const TX* t = tile + (size_t)(k & 1) * LS_RB * d;
    // ---- dot products of the block's rows with x
    double dp[LS_RB];
    if constexpr (FAST && sizeof(TX) == 4) {
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
            const float4 tv = *reinterpret_cast<const float4*>(
                reinterpret_cast<const float*>(t) + (size_t)r * d + j);
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
            double t0, t1, t2, t3;
            ld4(t + (size_t)r * d + j, t0, t1, t2, t3);
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
        if constexpr (FAST && sizeof(TX) == 4) {
          float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
          for (int r = 0; r < LS_RB; ++r)
            if (r < rows) {
              const float4 tv = *reinterpret_cast<const float4*>(
                  reinterpret_cast<const float*>(t) + (size_t)r * d + j);
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
              double t0, t1, t2, t3;
              ld4(t + (size_t)r * d + j, t0, t1, t2, t3);
              const double w = wr[r];
              acc[4 * c + 0] = fma(t0, w, acc[4 * c + 0]);
              acc[4 * c + 1] = fma(t1, w, acc[4 * c + 1]);
              acc[4 * c + 2] = fma(t2, w, acc[4 * c + 2]);
              acc[4 * c + 3] = fma(t3, w, acc[4 * c + 3]);
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


// ------------------------------------------------------------------------------- launch sink
// Where a sequence of kernels goes: straight onto a stream, or into a CUDA graph under
// construction as a chain of kernel nodes (each node depends on the previous one).  The same
// emit_* helpers below serve the plain entry point, the host-driven debug loop and the graph
// builder of the device-resident engine.
template <typename T> struct Ident { using type = T; };

struct Sink {
  cudaStream_t st = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphNode_t last = nullptr;
  cudaError_t err = cudaSuccess;
  int launches = 0;

  template <typename... KArgs>
  void launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
              typename Ident<KArgs>::type... args) {
    if (err != cudaSuccess) return;
    void* params[] = {(void*)&args...};
    ++launches;
    if (graph) {
      cudaKernelNodeParams kp = {};
      kp.func = (void*)kern;
      kp.gridDim = grid;
      kp.blockDim = block;
      kp.sharedMemBytes = (unsigned)smem;
      kp.kernelParams = params;
      cudaGraphNode_t node;
      err = cudaGraphAddKernelNode(&node, graph, last ? &last : nullptr, last ? 1 : 0, &kp);
      last = node;
    } else {
      err = cudaLaunchKernel((void*)kern, grid, block, params, smem, st);
    }
  }

  // WHILE node behind the last node; `body` receives the loop body graph
  void loop(cudaGraphConditionalHandle h, Sink* body) {
    if (err != cudaSuccess) return;
    cudaGraphNodeParams np = {};
    np.type = cudaGraphNodeTypeConditional;
    np.conditional.handle = h;
    np.conditional.type = cudaGraphCondTypeWhile;
    np.conditional.size = 1;
    cudaGraphNode_t node;
    err = cudaGraphAddNode(&node, graph, last ? &last : nullptr, last ? 1 : 0, &np);
    if (err != cudaSuccess) return;
    last = node;
    body->graph = np.conditional.phGraph_out[0];
    body->last = nullptr;
    body->err = cudaSuccess;
    body->launches = 0;
  }
};

using XtwxKernel = void (*)(const void*, const double*, const double*, long long, double*, double*,
                            double*, int, int, int, int);

struct XtwxPlan {
  XtwxKernel fn = nullptr;
  size_t smem = 0;
  int rb = 0;
};

// rows per streamed block: the largest whose two buffers fit next to x in shared memory
static int xtwx_rb(int d, int elt) {
  const int cands[4] = {8, 4, 2, 1};
  for (int i = 0; i < 4; ++i) {
    const int rb = cands[i];
    if ((size_t)2 * rb * d * elt + (size_t)(d + rb + rb * 8) * 8 <= (size_t)200 * 1024) return rb;
  }
  return 0;
}

template <typename TX>
static XtwxKernel xtwx_pick(int rb, bool fast) {
  if constexpr (sizeof(TX) == 4) {
    if (fast) {
      switch (rb) {
        case 8: return logistic_xtwx_kernel<float, 8, true>;
        case 4: return logistic_xtwx_kernel<float, 4, true>;
        case 2: return logistic_xtwx_kernel<float, 2, true>;
        default: return logistic_xtwx_kernel<float, 1, true>;
      }
    }
  }
  switch (rb) {
    case 8: return logistic_xtwx_kernel<TX, 8, false>;
    case 4: return logistic_xtwx_kernel<TX, 4, false>;
    case 2: return logistic_xtwx_kernel<TX, 2, false>;
    default: return logistic_xtwx_kernel<TX, 1, false>;
  }
}

// store: GADMM_STORE_F32 / GADMM_STORE_F64 shards.  hv == 2 (fp32 multiplies) exists for fp32
// shards only; with fp64 shards it silently means hv == 1 (exact).
static int xtwx_plan(int d, int store, int hv, XtwxPlan* out) {
  GADMM_REQUIRE(d > 0 && d % 4 == 0 && d <= LS_MAXC * LS_THREADS,
                "logistic_xtwx: d=%d must be a multiple of 4 and <= %d", d, LS_MAXC * LS_THREADS);
  const int elt = store == GADMM_STORE_F64 ? 8 : 4;
  const int rb = xtwx_rb(d, elt);
  GADMM_REQUIRE(rb > 0, "logistic_xtwx: d=%d does not fit in shared memory", d);
  out->rb = rb;
  out->smem = (size_t)2 * rb * d * elt + (size_t)(d + rb + rb * 8) * sizeof(double);
  const bool fast = hv == 2 && elt == 4;
  out->fn = elt == 8 ? xtwx_pick<double>(rb, false) : xtwx_pick<float>(rb, fast);
  return 0;
}

static int xtwx_configure() {
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (configured) return 0;
  const int cap = 220 * 1024;
  const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
  const int rbs[4] = {8, 4, 2, 1};
  for (int i = 0; i < 4; ++i) {
    GADMM_CUDA_OK(cudaFuncSetAttribute((const void*)xtwx_pick<float>(rbs[i], false), attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute((const void*)xtwx_pick<float>(rbs[i], true), attr, cap));
    GADMM_CUDA_OK(cudaFuncSetAttribute((const void*)xtwx_pick<double>(rbs[i], false), attr, cap));
  }
  configured = 1;
  return 0;
}

// One product v[n] = X_n' w(X_n x_n), n = 0..N-1, into `sink` (two kernels).
static void emit_xtwx(Sink& k, const XtwxPlan& pl, const void* X, const double* y, const double* x,
                      long long x_stride, double* Dw, double* v, int ldv, double* loss, int N,
                      int s, int d, int G, int hv, int shard_stride, double* scratch,
                      bool reduce = true) {
  double* part = scratch;
  double* losspart = scratch + (size_t)N * G * d;
  k.launch(pl.fn, dim3(G, N), dim3(LS_THREADS), pl.smem, X, y, x, x_stride, Dw, part, losspart, s,
           d, hv, shard_stride);
  if (!reduce) return;                 // the consumer kernel adds the G partial sums itself
  k.launch(logistic_reduce_kernel, dim3((d + 255) / 256, N), dim3(256), 0, (const double*)part,
           (const double*)losspart, v, loss, G, d, ldv, hv);
}

static int xtwx_entry(const void* X, int store, const double* y, const double* x, long long x_stride,
                      double* Dw, double* v, int ldv, double* loss, int N, int s, int d, int G,
                      int hv, int shard_stride, double* scratch, void* stream) {
  GADMM_REQUIRE(N > 0 && s > 0 && d > 0 && G > 0 && hv >= 0 && hv <= 2, "logistic_xtwx: bad shape");
  XtwxPlan pl;
  if (int e = xtwx_plan(d, store, hv, &pl)) return e;
  if (int e = xtwx_configure()) return e;
  Sink k;
  k.st = (cudaStream_t)stream;
  emit_xtwx(k, pl, X, y, x, x_stride, Dw, v, ldv, loss, N, s, d, G, hv, shard_stride, scratch);
  GADMM_CUDA_OK(k.err);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

// =============================================================================================
// Device-resident logistic GADMM (exact local solves by batched Newton-CG).
// Vector kernels: one CTA of 256 threads per row (= one worker of the colour being solved);
// per-row scalars live in global memory; "any row ..." decisions are taken by the last CTA
// (atomic ticket) which also sets the WHILE condition of the enclosing graph loop.
// =============================================================================================
constexpr int LG_CL = 4;      // CTAs (one thread-block cluster) per row in the CG / line-search steps
                              // when a GPU holds few rows; one CTA per row otherwise
constexpr int LG_THREADS = 1024;  // one CTA per row: wide CTAs keep the per-thread chains short when a GPU holds few rows
enum { C_OUTER = 0, C_NEWTON = 1, C_CG = 2, C_LS = 3 };

struct LogiConds {                 // conditional handles of the four loops (0 = host-driven mode)
  cudaGraphConditionalHandle h[4];
};

__host__ __device__ inline int logi_first(const GadmmLogi& p, int colour) {
  return (((colour - p.g0) % 2) + 2) % 2;
}
__host__ __device__ inline int logi_rows(const GadmmLogi& p, int colour) {
  const int first = logi_first(p, colour);
  return first < p.Nloc ? (p.Nloc - first + 1) / 2 : 0;
}

__host__ __device__ inline int logi_rmax(const GadmmLogi& p) {
  const int a = logi_rows(p, 0), b = logi_rows(p, 1);
  return a > b ? a : b;
}

__device__ __forceinline__ void logi_set_cond(const GadmmLogi& p, const LogiConds& lc, int which,
                                              unsigned int v) {
  p.cond[which] = v;
  if (lc.h[which] != 0) cudaGraphSetConditional(lc.h[which], v);
}

// A row of the batched solve is handled by a cluster of LG_CL CTAs, each owning a contiguous slice
// of the d columns: with few rows per GPU (32 at config 3 on 8 GPUs) one CTA per row would leave
// most SMs idle and the step would be bound by one SM's L2 bandwidth per row.
struct RowSlice {
  int n, j0, j1;
  unsigned rank;
};
template <int CL>
__device__ __forceinline__ RowSlice logi_slice(int d) {
  RowSlice r;
  r.n = blockIdx.x / CL;
  r.rank = blockIdx.x % CL;
  const int per = (d + CL - 1) / CL;
  r.j0 = (int)r.rank * per;
  r.j1 = min(d, r.j0 + per);
  return r;
}
// Sum of `v` over all threads of the cluster, identical in every thread (fixed order): block sum,
// then the LG_CL block sums are read through distributed shared memory.  `slot` = one shared
// double per call site (distinct call sites use distinct slots, so one cluster barrier each).
template <int CL>
__device__ __forceinline__ double cluster_sum(double v, double* red, double* slot) {
  v = block_sum(v, red);
  if constexpr (CL == 1) {
    return v;
  } else {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    if (threadIdx.x == 0) *slot = v;
    cl.sync();
    double t = 0.0;
#pragma unroll
    for (unsigned r = 0; r < (unsigned)CL; ++r) t += *cl.map_shared_rank(slot, r);
    return t;
  }
}
template <int CL>
__device__ __forceinline__ void cluster_exit() {   // peers may still be reading this CTA's slots
  if constexpr (CL > 1) cooperative_groups::this_cluster().sync();
}

// Ticket: returns true (to every thread) in the last CTA of the grid; `any` is OR-reduced over
// the grid through counters[slot+1].
__device__ __forceinline__ bool logi_last_cta(const GadmmLogi& p, int slot, bool any_here,
                                              unsigned int* s_ticket, bool* any_out) {
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (any_here) atomicAdd(p.counters + slot + 1, 1u);
    __threadfence();
    *s_ticket = atomicAdd(p.counters + slot, 1u);
  }
  __syncthreads();
  const bool last = *s_ticket == gridDim.x - 1;
  if (last) {
    __threadfence();
    *any_out = *reinterpret_cast<volatile unsigned int*>(p.counters + slot + 1) != 0u;
    __syncthreads();
    if (threadIdx.x == 0) {
      p.counters[slot] = 0u;
      p.counters[slot + 1] = 0u;
    }
  }
  return last;
}

// Start of the CG solve of H p = g for row n (zero start; a masked row keeps r = q = 0 and therefore
// p = 0), and the trial point / slope that go with p = 0.  Called by the kernels that produce g.
__device__ __forceinline__ void logi_cg_start(const GadmmLogi& p, int n, bool live, double s_gg) {
  const int tid = threadIdx.x, d = p.d;
  const double livef = live ? 1.0 : 0.0;
  for (int j = tid; j < d; j += LG_THREADS) {
    const size_t o = (size_t)n * d + j;
    const double rv = p.g[o] * livef;
    p.p[o] = 0.0;
    p.r[o] = rv;
    p.q[o] = rv;
    p.xtry[o] = p.x[o];
  }
  if (tid == 0) {
    const double gn = sqrt(s_gg);
    p.rs[n] = s_gg * livef;
    // Forcing term min(1e-2, |g|) |g|.  The local Hessians are mu I plus a small data part, so a
    // CG step shrinks the residual by ~50x: a tight forcing term costs one or two more Hessian
    // products per Newton step and saves whole Newton steps (each of which needs at least one
    // gradient pass, the more expensive of the two passes).
    p.eta[n] = fmin(gn, 1e-2) * gn;
    p.t[n] = livef;
    p.slope[n] = 0.0;
  }
}

// ---- start of a colour's half-iteration: neighbours' models -> linear term, warm start, and the
// gradient at the warm start from the data terms of the colour's last accepted evaluation (its
// models have not moved since, so no pass over the shards is needed).
//   local problem (group_ADMM_logistic.m:48-49, :83-85):
//   min_x loss_n(x) + mu/2 |x|^2 + lin'x,  mu = lam_log + c rho,
//   lin = -lam_{n-1} [n>0] + lam_n [n<N-1] - rho (theta_{n-1} + theta_{n+1})
__global__ void __launch_bounds__(LG_THREADS)
logi_setup_kernel(const GadmmLogi p, const LogiConds lc, const int colour) {
  __shared__ double red[32];
  __shared__ unsigned int s_ticket;
  const int n = blockIdx.x, tid = threadIdx.x, d = p.d;
  const int k = logi_first(p, colour) + 2 * n, g = p.g0 + k;
  const unsigned long long it = p.ctrl[0];
  const bool hasL = g > 0, hasR = g < p.N - 1;
  const unsigned long long need = colour == 0 ? it - 1ull : it;
  if (tid == 0 && p.nranks > 1 && ((hasL && k == 0) || (hasR && k == p.Nloc - 1))) {
    const unsigned long long t0 = globaltimer_ns();
    if (hasL && k == 0) wait_flag_ge(p.flags + g - 1, need);
    if (hasR && k == p.Nloc - 1) wait_flag_ge(p.flags + g + 1, need);
    atomicAdd(p.ctrl + 11, globaltimer_ns() - t0);          // time spent waiting for neighbours
  }
  __syncthreads();
  const double* tl = p.table + (size_t)(hasL ? g - 1 : g) * d;
  const double* tr = p.table + (size_t)(hasR ? g + 1 : g) * d;
  const double* ll = p.lam + (size_t)k * d;
  const double* lr = p.lam + (size_t)(k + 1) * d;
  const double mu = p.lam_log + ((hasL ? 1.0 : 0.0) + (hasR ? 1.0 : 0.0)) * p.rho;
  const int R = logi_rmax(p);
  const double* vl = p.vlast + ((size_t)colour * R + n) * d;
  double s_xx = 0.0, s_lx = 0.0, s_gg = 0.0;
  for (int j = tid; j < d; j += LG_THREADS) {
    double lin = 0.0, nb = 0.0;
    if (hasL) { lin -= ll[j]; nb += __ldcg(tl + j); }
    if (hasR) { lin += lr[j]; nb += __ldcg(tr + j); }
    lin -= p.rho * nb;
    const double xv = p.table[(size_t)g * d + j];
    const double gv = vl[j] + mu * xv + lin;
    p.lin[(size_t)n * d + j] = lin;
    p.x[(size_t)n * d + j] = xv;
    p.g[(size_t)n * d + j] = gv;
    s_xx = fma(xv, xv, s_xx);
    s_lx = fma(lin, xv, s_lx);
    s_gg = fma(gv, gv, s_gg);
  }
  s_xx = block_sum(s_xx, red);
  s_lx = block_sum(s_lx, red);
  s_gg = block_sum(s_gg, red);
  const double gn = sqrt(s_gg);
  const bool live = gn > p.tol * fmax(1.0, sqrt(s_xx));
  if (tid == 0) {
    p.mu[n] = mu;
    p.f[n] = p.losslast[(size_t)colour * R + n] + 0.5 * mu * s_xx + s_lx;
    p.gn[n] = gn;
    p.live[n] = live ? 1 : 0;
  }
  __syncthreads();                               // g, x of this row are in place
  logi_cg_start(p, n, live, s_gg);
  bool any;
  if (logi_last_cta(p, 0, live, &s_ticket, &any) && tid == 0) {
    p.ctrl[4] = 0ull;
    p.ctrl[5] = 0ull;
    logi_set_cond(p, lc, C_CG, 1u);              // a Newton step starts with at least one CG step
    logi_set_cond(p, lc, C_NEWTON, any ? 1u : 0u);
  }
}

// ---- one CG step.  The pass left G partial column sums per row in `scratch`; they are added here
// in a fixed order (deterministic) together with the regulariser mu q.  The step also keeps the
// line search's starting point up to date (xtry = x - p for live rows, slope = g'p), so that the
// line search can start as soon as the CG loop ends, whenever that is.
template <int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(LG_THREADS)
logi_cg_step_kernel(const GadmmLogi p, const LogiConds lc, const int colour) {
  __shared__ double red[32];
  __shared__ double slots[3];
  __shared__ unsigned int s_ticket;
  const int tid = threadIdx.x, d = p.d;
  const RowSlice rs_ = logi_slice<CL>(d);
  const int n = rs_.n;
  const int G = p.G[colour] > 0 ? p.G[colour] : 1;
  const double* part = p.scratch + (size_t)n * G * d;
  const double mu = p.mu[n], rs = p.rs[n], t = p.t[n], eta = p.eta[n];
  double qhq = 0.0;
  for (int j = rs_.j0 + tid; j < rs_.j1; j += LG_THREADS) {
    const size_t o = (size_t)n * d + j;
    const double qv = p.q[o];
    double hv = 0.0;
    for (int g = 0; g < G; ++g) hv += __ldcg(part + (size_t)g * d + j);
    hv = fma(mu, qv, hv);
    p.hq[o] = hv;
    qhq = fma(qv, hv, qhq);
  }
  qhq = cluster_sum<CL>(qhq, red, &slots[0]);
  const double alpha = qhq > 0.0 ? rs / qhq : 0.0;
  double rs_new = 0.0, slope = 0.0;
  for (int j = rs_.j0 + tid; j < rs_.j1; j += LG_THREADS) {
    const size_t o = (size_t)n * d + j;
    const double pv = fma(alpha, p.q[o], p.p[o]);
    p.p[o] = pv;
    const double rv = fma(-alpha, p.hq[o], p.r[o]);
    p.r[o] = rv;
    rs_new = fma(rv, rv, rs_new);
    slope = fma(p.g[o], pv, slope);
    p.xtry[o] = fma(-t, pv, p.x[o]);
  }
  rs_new = cluster_sum<CL>(rs_new, red, &slots[1]);
  slope = cluster_sum<CL>(slope, red, &slots[2]);
  const double beta = rs > 0.0 ? rs_new / rs : 0.0;
  for (int j = rs_.j0 + tid; j < rs_.j1; j += LG_THREADS) {
    const size_t o = (size_t)n * d + j;
    p.q[o] = fma(beta, p.q[o], p.r[o]);
  }
  if (tid == 0 && rs_.rank == 0) {
    p.rs[n] = rs_new;
    p.slope[n] = slope;
  }
  const bool more = sqrt(rs_new) > eta;
  bool any;
  const bool last = logi_last_cta(p, 2, more, &s_ticket, &any);
  if (last && tid == 0) {
    const unsigned long long c = p.ctrl[5] + 1ull;
    p.ctrl[5] = c;
    p.ctrl[6] = 0ull;
    p.ctrl[8] += 1ull;
    logi_set_cond(p, lc, C_LS, 1u);              // the line search evaluates at least once
    logi_set_cond(p, lc, C_CG, (any && c < (unsigned long long)p.max_cg) ? 1u : 0u);
  }
  cluster_exit<CL>();
}

// ---- Armijo test of the trial points (hq = X'w(X xtry), losstry from the pass); rows that fail
// halve their step and every row is evaluated again, so that the last evaluation is at the
// accepted points of ALL rows (its data terms and curvature state are reused afterwards).
template <int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(LG_THREADS)
logi_ls_step_kernel(const GadmmLogi p, const LogiConds lc, const int colour) {
  __shared__ double red[32];
  __shared__ double slots[3];
  __shared__ unsigned int s_ticket;
  __shared__ double s_loss;
  const int tid = threadIdx.x, d = p.d, rows = gridDim.x / CL;
  const RowSlice rs_ = logi_slice<CL>(d);
  const int n = rs_.n;
  const int G = p.G[colour] > 0 ? p.G[colour] : 1;
  const double* part = p.scratch + (size_t)n * G * d;
  // per-row scalars are read before the first cluster barrier: rank 0 rewrites some of them below
  const double mu = p.mu[n], f = p.f[n], t = p.t[n], slope0 = p.slope[n];
  const bool live = p.live[n] != 0;
  const bool capped = p.ctrl[6] + 1ull >= (unsigned long long)p.max_ls;
  if (tid == 0) {                                // loss of the trial point: G partials, fixed order
    const double* lp = p.scratch + (size_t)rows * G * d + (size_t)n * G;
    double l = 0.0;
    for (int g = 0; g < G; ++g) l += __ldcg(lp + g);
    if (rs_.rank == 0) p.losstry[n] = l;
    s_loss = l;
  }
  double s_xx = 0.0, s_lx = 0.0, s_alx = 0.0;
  for (int j = rs_.j0 + tid; j < rs_.j1; j += LG_THREADS) {
    const size_t o = (size_t)n * d + j;
    const double xv = p.xtry[o];
    double v = 0.0;                              // data part of the trial gradient
    for (int g = 0; g < G; ++g) v += __ldcg(part + (size_t)g * d + j);
    p.hq[o] = v;
    const double lx = p.lin[o] * xv;
    s_xx = fma(xv, xv, s_xx);
    s_lx += lx;
    s_alx += fabs(lx);
  }
  s_xx = cluster_sum<CL>(s_xx, red, &slots[0]);
  s_lx = cluster_sum<CL>(s_lx, red, &slots[1]);
  s_alx = cluster_sum<CL>(s_alx, red, &slots[2]);
  const double loss_try = s_loss;                // (the barriers above make it visible)
  const double f_try = loss_try + 0.5 * mu * s_xx + s_lx;
  // A predicted decrease below the rounding noise of f cannot be tested: accept it.  f is the sum
  // of three terms that may cancel (loss, quadratic, linear), each accumulated over thousands of
  // products, so the noise is taken relative to their magnitudes, not to |f|.  Without this,
  // rows that are converged to rounding level burn line-search trials (a full pass over the
  // shards each, for every row of the colour) on coin-flip Armijo tests.
  const double noise = 256.0 * 2.220446049250313e-16 *
                       (fabs(loss_try) + 0.5 * mu * s_xx + s_alx + 1.0);
  const bool bad = live && t > 1e-9 && (f_try > f - 1e-4 * t * slope0 + noise) && !capped;
  if (bad) {
    const double t2 = 0.5 * t;
    for (int j = rs_.j0 + tid; j < rs_.j1; j += LG_THREADS) {
      const size_t o = (size_t)n * d + j;
      p.xtry[o] = fma(-t2, p.p[o], p.x[o]);
    }
    if (tid == 0 && rs_.rank == 0) p.t[n] = t2;
  }
  // f of the evaluated point; for a failing row it is overwritten by the next trial
  if (tid == 0 && rs_.rank == 0) p.eta[n] = f_try;   // eta is free during the line search: holds f_try
  bool any;
  const bool last = logi_last_cta(p, 4, bad, &s_ticket, &any);
  if (last && tid == 0) {
    p.ctrl[6] += 1ull;
    p.ctrl[9] += 1ull;
    logi_set_cond(p, lc, C_LS, any ? 1u : 0u);
  }
  cluster_exit<CL>();
}

// ---- accept the step: x, f, g at the new point; Newton loop condition
__global__ void __launch_bounds__(LG_THREADS)
logi_newton_next_kernel(const GadmmLogi p, const LogiConds lc, const int colour) {
  __shared__ double red[32];
  __shared__ unsigned int s_ticket;
  const int n = blockIdx.x, tid = threadIdx.x, d = p.d;
  const double mu = p.mu[n];
  const int R = logi_rmax(p);
  double* vl = p.vlast + ((size_t)colour * R + n) * d;
  double s_xx = 0.0, s_gg = 0.0;
  for (int j = tid; j < d; j += LG_THREADS) {
    const size_t o = (size_t)n * d + j;
    const double xv = p.xtry[o], v = p.hq[o];
    const double gv = v + mu * xv + p.lin[o];
    p.x[o] = xv;
    p.g[o] = gv;
    vl[j] = v;
    s_xx = fma(xv, xv, s_xx);
    s_gg = fma(gv, gv, s_gg);
  }
  s_xx = block_sum(s_xx, red);
  s_gg = block_sum(s_gg, red);
  const double gn = sqrt(s_gg);
  const bool live = gn > p.tol * fmax(1.0, sqrt(s_xx));
  if (tid == 0) {
    p.f[n] = p.eta[n];
    p.gn[n] = gn;
    p.live[n] = live ? 1 : 0;
    p.losslast[(size_t)colour * R + n] = p.losstry[n];
  }
  __syncthreads();
  logi_cg_start(p, n, live, s_gg);               // ready for the next Newton step, if there is one
  bool any;
  if (logi_last_cta(p, 0, live, &s_ticket, &any) && tid == 0) {
    const unsigned long long c = p.ctrl[4] + 1ull;
    p.ctrl[4] = c;
    p.ctrl[5] = 0ull;
    p.ctrl[10] += 1ull;
    logi_set_cond(p, lc, C_CG, 1u);
    logi_set_cond(p, lc, C_NEWTON, (any && c < (unsigned long long)p.max_newton) ? 1u : 0u);
  }
}

// ---- end of a colour's half-iteration: models into the table (and, for the two boundary
// workers, straight into the neighbouring GPU's table over NVLink), then the version flags
__global__ void __launch_bounds__(LG_THREADS)
logi_publish_kernel(const GadmmLogi p, const int colour) {
  const int n = blockIdx.x, tid = threadIdx.x, d = p.d;
  const int k = logi_first(p, colour) + 2 * n, g = p.g0 + k;
  const unsigned long long it = p.ctrl[0];
  double* pl = (p.nranks > 1 && k == 0 && g > 0) ? p.table_peer[p.rank - 1] + (size_t)g * d : nullptr;
  double* pr = (p.nranks > 1 && k == p.Nloc - 1 && g < p.N - 1)
                   ? p.table_peer[p.rank + 1] + (size_t)g * d : nullptr;
  for (int j = tid; j < d; j += LG_THREADS) {
    const double xv = p.x[(size_t)n * d + j];
    p.table[(size_t)g * d + j] = xv;
    if (pl) pl[j] = xv;
    if (pr) pr[j] = xv;
  }
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    if (pl) st_release_sys(p.flags_peer[p.rank - 1] + g, it);
    if (pr) st_release_sys(p.flags_peer[p.rank + 1] + g, it);
    st_release_sys(p.flags + g, it);
  }
}

// ---- end of an iteration: edge duals lam_e += rho (theta_e - theta_{e+1}) (group_ADMM_logistic.m
// :90-91), objective sum_n lam_log/2 |theta_n|^2 + loss_n(theta_n) (:96-103) from the losses of the
// two half-iterations' last evaluations, all-rank sum through peer stores + flags, stop test
// (:104-110), clock, iteration counter, outer loop condition.  grid = Nloc + 1 CTAs:
// CTA k handles edge (g0-1+k, g0+k) and, for k < Nloc, worker k's objective piece.
__global__ void __launch_bounds__(LG_THREADS)
logi_dual_obj_kernel(const GadmmLogi p, const LogiConds lc) {
  __shared__ double red[32];
  __shared__ unsigned int s_ticket;
  const int k = blockIdx.x, tid = threadIdx.x, d = p.d;
  const unsigned long long it = p.ctrl[0];
  const int gl = p.g0 - 1 + k, gr = p.g0 + k;              // the edge's end points
  const bool edge = gl >= 0 && gr <= p.N - 1;
  if (tid == 0 && p.nranks > 1 && edge && (k == 0 || k == p.Nloc)) {
    const unsigned long long t0 = globaltimer_ns();
    if (k == 0) wait_flag_ge(p.flags + gl, it);            // left neighbour GPU's last worker
    if (k == p.Nloc) wait_flag_ge(p.flags + gr, it);       // right neighbour GPU's first worker
    atomicAdd(p.ctrl + 12, globaltimer_ns() - t0);
  }
  __syncthreads();
  double s_xx = 0.0;
  for (int j = tid; j < d; j += LG_THREADS) {
    if (edge) {
      const double a = __ldcg(p.table + (size_t)gl * d + j), b = __ldcg(p.table + (size_t)gr * d + j);
      p.lam[(size_t)k * d + j] += p.rho * (a - b);
    }
    if (k < p.Nloc) {
      const double xv = __ldcg(p.table + (size_t)gr * d + j);
      s_xx = fma(xv, xv, s_xx);
    }
  }
  s_xx = block_sum(s_xx, red);
  if (tid == 0 && k < p.Nloc) {
    const int R = logi_rmax(p);
    const int colour = gr & 1, n = (k - logi_first(p, colour)) / 2;
    p.objw[k] = p.losslast[(size_t)colour * R + n] + 0.5 * p.lam_log * s_xx;
  }
  bool any;
  if (!logi_last_cta(p, 6, false, &s_ticket, &any)) return;
  double part = 0.0;
  for (int i = tid; i < p.Nloc; i += LG_THREADS) part += __ldcg(p.objw + i);
  part = block_sum(part, red);
  if (tid >= 32) return;
  const int lane = tid;
  double total = part;
  if (p.nranks > 1) {
    const int ring = (int)(it & 3ull) * p.nranks;
    double v = 0.0;
    if (lane < p.nranks) {
      *reinterpret_cast<volatile double*>(p.objpart_peer[lane] + ring + p.rank) = part;
      __threadfence_system();
      st_release_sys(p.objflag_peer[lane] + ring + p.rank, it);
      const unsigned long long t0 = globaltimer_ns();
      wait_flag_ge(p.objflag + ring + lane, it);
      v = *reinterpret_cast<volatile double*>(p.objpart + ring + lane);
      if (lane != p.rank) atomicMax(p.ctrl + 14, globaltimer_ns() - t0);
    }
    total = 0.0;
    for (int q = 0; q < p.nranks; ++q) total += __shfl_sync(0xffffffffu, v, q);
    if (lane == 0) { p.ctrl[13] += p.ctrl[14]; p.ctrl[14] = 0ull; }
  }
  if (lane == 0) {
    const double loss = fabs(total - p.obj0);
    p.obj_hist[it] = total;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    bool hit = false;
    if (loss < p.acc && p.ctrl[2] != 0ull) {
      if (p.ctrl[1] == 0ull) p.ctrl[1] = it;
      hit = true;
    }
    p.ctrl[0] = it + 1ull;
    const bool more = !hit && it + 1ull <= p.ctrl[3] && it + 1ull <= (unsigned long long)p.max_iter;
    logi_set_cond(p, lc, C_OUTER, more ? 1u : 0u);
  }
}

// ---- entry of a launch: decide whether there is anything to do (resume after a stop / cap)
__global__ void logi_begin_kernel(const GadmmLogi p, const LogiConds lc) {
  const unsigned long long it = p.ctrl[0];
  const bool stopped = p.ctrl[2] != 0ull && p.ctrl[1] != 0ull;
  const bool more = !stopped && it <= p.ctrl[3] && it <= (unsigned long long)p.max_iter;
  if (p.t_hist[0] == 0ull) p.t_hist[0] = globaltimer_ns();   // clock origin: the first launch
  logi_set_cond(p, lc, C_OUTER, more ? 1u : 0u);
}

// ---- conditional-node probe: body increments a counter until it reaches n
__global__ void loop_probe_kernel(cudaGraphConditionalHandle h, unsigned int* counter, unsigned int n) {
  const unsigned int c = *counter + 1u;
  *counter = c;
  cudaGraphSetConditional(h, c < n ? 1u : 0u);
}

// few rows: a cluster of LG_CL CTAs per row keeps the SMs busy; many rows: one CTA per row
static bool logi_wide(int rows) { return rows * LG_CL <= 148; }   // one wave of 1024-thread CTAs

// -------------------------------------------------------------------------- sequence emitters
struct LogiPlans {
  XtwxPlan grad, hv;
};

static int logi_plans(const GadmmLogi& p, LogiPlans* pl) {
  GADMM_REQUIRE(p.N > 0 && p.Nloc > 0 && p.s > 0 && p.nranks >= 1 && p.nranks <= 32,
                "logi: bad problem description");
  GADMM_REQUIRE(p.max_newton > 0 && p.max_cg > 0 && p.max_ls > 0 && p.max_iter > 0,
                "logi: iteration caps must be positive");
  if (int e = xtwx_plan(p.d, p.store, 0, &pl->grad)) return e;
  if (int e = xtwx_plan(p.d, p.store, p.hv_mode == 2 ? 2 : 1, &pl->hv)) return e;
  return xtwx_configure();
}

// shard pointers of a colour (rows n = 0.. <-> local workers first, first+2, ...)
struct LogiShards {
  const void* X;
  const double* y;
  double* Dw;
  int rows, G;
};
static LogiShards logi_shards(const GadmmLogi& p, int colour) {
  const int first = logi_first(p, colour);
  const size_t esz = p.store == GADMM_STORE_F64 ? 8 : 4;
  LogiShards sh;
  sh.X = reinterpret_cast<const unsigned char*>(p.X) + (size_t)first * p.s * p.d * esz;
  sh.y = p.y + (size_t)first * p.s;
  sh.Dw = p.Dw + (size_t)first * p.s;
  sh.rows = logi_rows(p, colour);
  sh.G = p.G[colour] > 0 ? p.G[colour] : 1;
  return sh;
}

// reduce = false: the G partial sums stay in p.scratch for the consumer kernel (cg_step / ls_step)
static void emit_eval(Sink& k, const GadmmLogi& p, const LogiPlans& pl, int colour, const double* x,
                      long long x_stride, double* v, double* loss, bool reduce) {
  const LogiShards sh = logi_shards(p, colour);
  emit_xtwx(k, pl.grad, sh.X, sh.y, x, x_stride, sh.Dw, v, p.d, loss, sh.rows, p.s, p.d, sh.G, 0, 2,
            p.scratch, reduce);
}
static void emit_hessvec(Sink& k, const GadmmLogi& p, const LogiPlans& pl, int colour) {
  const LogiShards sh = logi_shards(p, colour);
  emit_xtwx(k, pl.hv, sh.X, sh.y, p.q, p.d, sh.Dw, p.hq, p.d, p.losstry, sh.rows, p.s, p.d, sh.G,
            p.hv_mode == 2 ? 2 : 1, 2, p.scratch, false);
}

// One colour's half-iteration as a graph: setup -> WHILE newton { WHILE cg { X'DXq, cg_step } ->
// WHILE ls { evaluate(xtry), ls_step } -> newton_next } -> publish
static void graph_half(Sink& k, const GadmmLogi& p, const LogiPlans& pl, const LogiConds& lc,
                       int colour) {
  const int rows = logi_rows(p, colour);
  if (rows == 0) return;
  const dim3 grid(rows), block(LG_THREADS);
  k.launch(logi_setup_kernel, grid, block, 0, p, lc, colour);
  Sink nw;
  k.loop(lc.h[C_NEWTON], &nw);
  if (k.err != cudaSuccess) return;
  {
    Sink cg;
    nw.loop(lc.h[C_CG], &cg);
    if (nw.err == cudaSuccess) {
      emit_hessvec(cg, p, pl, colour);
      if (logi_wide(rows)) cg.launch(logi_cg_step_kernel<LG_CL>, dim3(rows * LG_CL), block, 0, p, lc, colour);
      else cg.launch(logi_cg_step_kernel<1>, grid, block, 0, p, lc, colour);
      if (cg.err != cudaSuccess) nw.err = cg.err;
    }
    Sink ls;
    nw.loop(lc.h[C_LS], &ls);
    if (nw.err == cudaSuccess) {
      emit_eval(ls, p, pl, colour, p.xtry, p.d, p.hq, p.losstry, false);
      if (logi_wide(rows)) ls.launch(logi_ls_step_kernel<LG_CL>, dim3(rows * LG_CL), block, 0, p, lc, colour);
      else ls.launch(logi_ls_step_kernel<1>, grid, block, 0, p, lc, colour);
      if (ls.err != cudaSuccess) nw.err = ls.err;
    }
    nw.launch(logi_newton_next_kernel, grid, block, 0, p, lc, colour);
    if (nw.err != cudaSuccess) k.err = nw.err;
  }
  k.launch(logi_publish_kernel, grid, block, 0, p, colour);
}

struct LogiGraph {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
};

static int read_cond(const GadmmLogi& p, int which, cudaStream_t st, unsigned int* out) {
  GADMM_CUDA_OK(cudaMemcpyAsync(out, p.cond + which, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
  GADMM_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
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
  return xtwx_entry(X, GADMM_STORE_F32, y, x, x_stride, Dw, v, ldv, loss, N, s, d, G, hv,
                    shard_stride, scratch, stream);
}

int gadmm_logistic_xtwx_ex(const void* X, int store, const double* y, const double* x,
                           long long x_stride, double* Dw, double* v, int ldv, double* loss, int N,
                           int s, int d, int G, int hv, int shard_stride, double* scratch,
                           void* stream) {
  GADMM_REQUIRE(store == GADMM_STORE_F64 || store == GADMM_STORE_F32, "logistic_xtwx: bad store %d",
                store);
  return xtwx_entry(X, store, y, x, x_stride, Dw, v, ldv, loss, N, s, d, G, hv, shard_stride,
                    scratch, stream);
}

int gadmm_logi_rows(const GadmmLogi* p, int colour) { return p ? logi_rows(*p, colour & 1) : 0; }

int gadmm_logi_prime(const GadmmLogi* p, void* stream) {
  GADMM_REQUIRE(p != nullptr, "logi: null descriptor");
  LogiPlans pl;
  if (int e = logi_plans(*p, &pl)) return e;
  Sink k;
  k.st = (cudaStream_t)stream;
  const int R = logi_rmax(*p);
  for (int colour = 0; colour < 2; ++colour) {
    if (logi_rows(*p, colour) == 0) continue;
    const int first = logi_first(*p, colour);
    emit_eval(k, *p, pl, colour, p->table + (size_t)(p->g0 + first) * p->d, 2ll * p->d,
              p->vlast + (size_t)colour * R * p->d, p->losslast + (size_t)colour * R, true);
  }
  GADMM_CUDA_OK(k.err);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_logi_graph_create(const GadmmLogi* p, void** handle_out) {
  GADMM_REQUIRE(p != nullptr && handle_out != nullptr, "logi: null argument");
  LogiPlans pl;
  if (int e = logi_plans(*p, &pl)) return e;
  LogiGraph* lg = new LogiGraph();
  const char* stage = "cudaGraphCreate";
  cudaError_t err = cudaGraphCreate(&lg->graph, 0);
  // one handle per WHILE node: the iteration loop, and (Newton, CG, line search) per colour
  LogiConds lc[2] = {};
  cudaGraphConditionalHandle outer = 0;
  if (err == cudaSuccess) {
    stage = "cudaGraphConditionalHandleCreate";
    err = cudaGraphConditionalHandleCreate(&outer, lg->graph, 0, 0);
  }
  for (int c = 0; c < 2 && err == cudaSuccess; ++c) {
    lc[c].h[C_OUTER] = outer;
    if (logi_rows(*p, c) == 0) continue;
    for (int i = 1; i < 4 && err == cudaSuccess; ++i)
      err = cudaGraphConditionalHandleCreate(&lc[c].h[i], lg->graph, 0, 0);
  }
  if (err == cudaSuccess) {
    stage = "adding nodes";
    Sink root;
    root.graph = lg->graph;
    root.launch(logi_begin_kernel, dim3(1), dim3(1), 0, *p, lc[0]);
    Sink body;
    root.loop(outer, &body);
    if (root.err == cudaSuccess) {
      graph_half(body, *p, pl, lc[0], 0);
      graph_half(body, *p, pl, lc[1], 1);
      body.launch(logi_dual_obj_kernel, dim3(p->Nloc + 1), dim3(LG_THREADS), 0, *p, lc[0]);
      if (body.err != cudaSuccess) root.err = body.err;
    }
    err = root.err;
  }
  if (err == cudaSuccess) stage = "cudaGraphInstantiate";
  if (err == cudaSuccess) err = cudaGraphInstantiate(&lg->exec, lg->graph, 0);
  if (err != cudaSuccess) {
    if (lg->graph) cudaGraphDestroy(lg->graph);
    delete lg;
    cudaGetLastError();
    return set_error((int)err, "logi: building the conditional graph failed in %s: %s", stage,
                     cudaGetErrorString(err));
  }
  *handle_out = lg;
  return 0;
}

int gadmm_logi_graph_launch(void* handle, void* stream) {
  GADMM_REQUIRE(handle != nullptr, "logi: null graph handle");
  GADMM_CUDA_OK(cudaGraphLaunch(static_cast<LogiGraph*>(handle)->exec, (cudaStream_t)stream));
  return 0;
}

int gadmm_logi_graph_destroy(void* handle) {
  if (!handle) return 0;
  LogiGraph* lg = static_cast<LogiGraph*>(handle);
  if (lg->exec) cudaGraphExecDestroy(lg->exec);
  if (lg->graph) cudaGraphDestroy(lg->graph);
  delete lg;
  return 0;
}

/* Host-driven twin of the graph: the same kernels in the same order, every loop test read back
 * from p->cond (one stream synchronisation each). */
int gadmm_logi_run_host(const GadmmLogi* p, void* stream) {
  GADMM_REQUIRE(p != nullptr, "logi: null descriptor");
  LogiPlans pl;
  if (int e = logi_plans(*p, &pl)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  const LogiConds lc = {};
  Sink k;
  k.st = st;
  unsigned int c = 0;
  k.launch(logi_begin_kernel, dim3(1), dim3(1), 0, *p, lc);
  GADMM_CUDA_OK(k.err);
  if (int e = read_cond(*p, C_OUTER, st, &c)) return e;
  while (c) {
    for (int colour = 0; colour < 2; ++colour) {
      const int rows = logi_rows(*p, colour);
      if (rows == 0) continue;
      const dim3 grid(rows), block(LG_THREADS);
      k.launch(logi_setup_kernel, grid, block, 0, *p, lc, colour);
      GADMM_CUDA_OK(k.err);
      unsigned int cn = 0;
      if (int e = read_cond(*p, C_NEWTON, st, &cn)) return e;
      while (cn) {
        unsigned int cc = 1;
        while (cc) {
          emit_hessvec(k, *p, pl, colour);
          if (logi_wide(rows)) k.launch(logi_cg_step_kernel<LG_CL>, dim3(rows * LG_CL), block, 0, *p, lc, colour);
          else k.launch(logi_cg_step_kernel<1>, grid, block, 0, *p, lc, colour);
          GADMM_CUDA_OK(k.err);
          if (int e = read_cond(*p, C_CG, st, &cc)) return e;
        }
        unsigned int cl = 1;
        while (cl) {
          emit_eval(k, *p, pl, colour, p->xtry, p->d, p->hq, p->losstry, false);
          if (logi_wide(rows)) k.launch(logi_ls_step_kernel<LG_CL>, dim3(rows * LG_CL), block, 0, *p, lc, colour);
          else k.launch(logi_ls_step_kernel<1>, grid, block, 0, *p, lc, colour);
          GADMM_CUDA_OK(k.err);
          if (int e = read_cond(*p, C_LS, st, &cl)) return e;
        }
        k.launch(logi_newton_next_kernel, grid, block, 0, *p, lc, colour);
        GADMM_CUDA_OK(k.err);
        if (int e = read_cond(*p, C_NEWTON, st, &cn)) return e;
      }
      k.launch(logi_publish_kernel, grid, block, 0, *p, colour);
    }
    k.launch(logi_dual_obj_kernel, dim3(p->Nloc + 1), dim3(LG_THREADS), 0, *p, lc);
    GADMM_CUDA_OK(k.err);
    if (int e = read_cond(*p, C_OUTER, st, &c)) return e;
  }
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

/* Conditional-node probe: a WHILE loop whose body increments a device counter until it is n.
 * Returns 0 when the loop ran exactly n times. */
int gadmm_graph_loops_selftest(int n, void* stream) {
  GADMM_REQUIRE(n >= 1, "selftest: n must be >= 1");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* counter = nullptr;
  GADMM_CUDA_OK(cudaMalloc(&counter, sizeof(unsigned int)));
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaError_t err = cudaMemsetAsync(counter, 0, sizeof(unsigned int), st);
  if (err == cudaSuccess) err = cudaGraphCreate(&graph, 0);
  cudaGraphConditionalHandle h = 0;
  if (err == cudaSuccess) err = cudaGraphConditionalHandleCreate(&h, graph, 1, cudaGraphCondAssignDefault);
  if (err == cudaSuccess) {
    Sink root, body;
    root.graph = graph;
    root.loop(h, &body);
    if (root.err == cudaSuccess) {
      body.launch(loop_probe_kernel, dim3(1), dim3(1), 0, h, counter, (unsigned int)n);
      root.err = body.err;
    }
    err = root.err;
  }
  if (err == cudaSuccess) err = cudaGraphInstantiate(&exec, graph, 0);
  if (err == cudaSuccess) err = cudaGraphLaunch(exec, st);
  unsigned int got = 0;
  if (err == cudaSuccess) err = cudaMemcpyAsync(&got, counter, sizeof(got), cudaMemcpyDeviceToHost, st);
  if (err == cudaSuccess) err = cudaStreamSynchronize(st);
  if (exec) cudaGraphExecDestroy(exec);
  if (graph) cudaGraphDestroy(graph);
  cudaFree(counter);
  if (err != cudaSuccess) {
    cudaGetLastError();
    return set_error((int)err, "graph loop selftest: %s", cudaGetErrorString(err));
  }
  GADMM_REQUIRE(got == (unsigned int)n, "graph loop selftest: body ran %u times, expected %d", got, n);
  return 0;
}

}  // extern "C"
