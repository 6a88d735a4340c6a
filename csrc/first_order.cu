// First-order baselines on the GPU (K7 + K8 of SURVEY.md section 2.4):
//   GD, LAG-PS, LAG-WK, cyclic IAG, randomised IAG  (shared iterate, "parameter server")
//   DGD, dual averaging                              (one iterate per worker, chain stencil)
// for least squares and l2-regularised logistic regression.
//
// An iteration is  [gradient provider]  +  [one fused update kernel]:
//   provider (linear)   packed-symmetric SYMV  G_n theta  with the epilogue  - b_n  and the
//                       per-worker objective 0.5 th'G th - b'th + 0.5 y'y   (symv_grad_kernel)
//   provider (logistic) -X_n'(y ./ (1+exp(y .* X_n th))) + lam th  and the per-worker loss
//   update              per-worker trigger tests / stale-gradient table / stencil, then the last
//                       CTA (atomic ticket) does the server sum, the step, the objective, the
//                       stop test and advances the device-resident iteration counter, so that a
//                       window of iterations replays as one CUDA graph.
// Reference semantics: GD_DGD_LAG.m:86-370, GD_DGD_LAG_logistic.m:89-392, dual_averaging.m:17-75,
// dual_averaging_logisticReg.m:17-80 (quirks: initial gradients of ones, iteration-1 guards,
// worker-1 refresh in LAG-PS, idle +1 of comm_iter every 1000 iterations).
#include "../include/gadmm_b200.h"
#include "symv_core.cuh"

namespace gadmm {

// ------------------------------------------------------------------ provider: linear gradient
struct StridedX {
  const double* x;
  __device__ __forceinline__ double operator()(int j) const { return x[j]; }
};

template <typename T>
__global__ void __launch_bounds__(kSymvThreads, 2)
symv_grad_kernel(const void* const* __restrict__ Sptr, const double* __restrict__ x,
                 long long x_stride, const double* __restrict__ b, const double* __restrict__ hyy,
                 double* __restrict__ Y, double* __restrict__ fw, int dp,
                 double* __restrict__ scratch, unsigned int* __restrict__ counters) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SymvSmem<T>& sm = *reinterpret_cast<SymvSmem<T>*>(smem_raw);
  const int nT = dp / kT, npairs = (nT + 1) / 2;
  const int k = blockIdx.x / npairs, p = blockIdx.x % npairs;
  const size_t per = (size_t)dp + colpart_elems(dp);
  double* rowpart = scratch + (size_t)k * per;
  double* colpart = rowpart + dp;
  StridedX xf{x + (size_t)k * x_stride};
  const bool last = symv_cta<T>(sm, reinterpret_cast<const T*>(Sptr[k]), nT, p, xf, rowpart,
                                colpart, counters + k, npairs);
  if (!last) return;
  const double* bk = b ? b + (size_t)k * dp : nullptr;
  double s1 = 0.0, s2 = 0.0;
  for (int j = threadIdx.x; j < dp; j += blockDim.x) {
    const double gx = symv_gather(rowpart, colpart, nT, j);
    const double xj = xf(j), bj = bk ? bk[j] : 0.0;
    Y[(size_t)k * dp + j] = gx - bj;
    s1 = fma(xj, gx, s1);
    s2 = fma(bj, xj, s2);
  }
  if (fw) {
    double* red = &sm.colbuf[0][0][0];
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    if (threadIdx.x == 0) fw[k] = 0.5 * s1 - s2 + (hyy ? hyy[k] : 0.0);
  }
}

// ---------------------------------------------------------------- provider: logistic gradient
// One CTA per worker.  X_n is s x d row-major fp64; warp-per-row for z = X th (coalesced along
// features), thread-per-feature for X'w (coalesced along features as well).
//   Y[n]  = -X_n'(y ./ (1 + exp(y .* z))) + lam*th        (logReg_GD.m:9 without the duals)
//   fw[n] = lam/2 ||th||^2 + sum log(1 + exp(-y .* z))      (group_ADMM_logistic_GD.m:114)
//   wout  = sigma(1-sigma) per sample (Newton weights), optional
__global__ void __launch_bounds__(256)
logistic_grad_kernel(const double* __restrict__ X, const double* __restrict__ y,
                     const double* __restrict__ th, long long th_stride, double lam, int s, int d,
                     int dp, double* __restrict__ Y, double* __restrict__ fw,
                     double* __restrict__ wout, double* __restrict__ wbuf) {
  __shared__ double red[32];
  const int n = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double* Xn = X + (size_t)n * s * d;
  const double* yn = y + (size_t)n * s;
  const double* t = th + (size_t)n * th_stride;
  double* w = wbuf + (size_t)n * s;            // per-sample weights y/(1+exp(y z)) (global scratch)
  double floss = 0.0;
  for (int r = warp; r < s; r += 8) {
    double z = 0.0;
    for (int j = lane; j < d; j += 32) z = fma(Xn[(size_t)r * d + j], t[j], z);
    z = warp_sum(z);
    if (lane == 0) {
      const double yz = yn[r] * z;
      const double e = exp(yz);               // naive forms, as the reference (quirk 8)
      w[r] = yn[r] / (1.0 + e);
      floss += log(1.0 + exp(-yz));
      if (wout) { const double sg = 1.0 / (1.0 + e); wout[(size_t)n * s + r] = sg * (1.0 - sg); }
    }
  }
  __syncthreads();
  double nrm = 0.0;
  for (int j = tid; j < dp; j += 256) {
    double g = 0.0;
    if (j < d) {
      for (int r = 0; r < s; ++r) g = fma(Xn[(size_t)r * d + j], w[r], g);
      g = -g + lam * t[j];
      nrm = fma(t[j], t[j], nrm);
    }
    Y[(size_t)n * dp + j] = g;
  }
  if (fw) {
    floss = block_sum(floss, red);
    nrm = block_sum(nrm, red);
    if (tid == 0) fw[n] = 0.5 * lam * nrm + floss;
  }
}

}  // namespace gadmm

namespace gadmm {

constexpr int kSlots = 10;   // triggerslot (GD_DGD_LAG.m:18)

// ---------------------------------------------------------- shared-iterate update (K7)
// grid = N CTAs x 256 threads.  CTA n: worker n's upload decision.  Last CTA: server.
__global__ void __launch_bounds__(256)
shared_update_kernel(const GadmmShared p) {
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  __shared__ int s_up;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int n = blockIdx.x, tid = threadIdx.x, dp = p.dp;
  const int gn = p.gid0 + n;                  // global worker id (workers are split over the GPUs)
  const double* fresh = p.Y + (size_t)n * dp;
  double* gh = p.ghat + (size_t)n * dp;

  if (p.kind == GADMM_FO_LAG_PS || p.kind == GADMM_FO_LAG_WK) {
    if (p.kind == GADMM_FO_LAG_PS && gn == 0 && it > 1ull)       // GD_DGD_LAG.m:204-209
      for (int j = tid; j < dp; j += 256) gh[j] = fresh[j];
    if (it > (unsigned long long)kSlots) {
      double trig = 0.0;
      for (int q = 0; q < kSlots; ++q) trig += p.dnorm[q];
      double v = 0.0;
      if (p.kind == GADMM_FO_LAG_PS) {
        const double* thh = p.theta_hat + (size_t)n * dp;
        for (int j = tid; j < p.d; j += 256) { const double e = thh[j] - p.theta[j]; v = fma(e, e, v); }
        v = block_sum(v, red) * p.Hmax[n] * p.Hmax[n];          // :219
      } else {
        for (int j = tid; j < p.d; j += 256) { const double e = fresh[j] - gh[j]; v = fma(e, e, v); }
        v = block_sum(v, red);                                  // :292
      }
      if (tid == 0) s_up = (v > p.thr * trig) ? 1 : 0;
      __syncthreads();
      if (s_up) {
        for (int j = tid; j < dp; j += 256) {
          gh[j] = fresh[j];
          if (p.kind == GADMM_FO_LAG_PS) p.theta_hat[(size_t)n * dp + j] = p.theta[j];
        }
        if (tid == 0) atomicAdd(p.counters + 1, 1u);          // uploads of this iteration
      }
    }
  } else if (p.kind == GADMM_FO_IAG_CYCLIC) {
    if (it > 1ull && gn == (int)(it % (unsigned long long)p.Nglob))   // i=mod(iter,N)+1 (:337)
      for (int j = tid; j < dp; j += 256) gh[j] = fresh[j];
  } else if (p.kind == GADMM_FO_IAG_RANDOM) {
    const double u = p.uni[it];
    int i = 0;
    while (i < p.Nglob - 1 && !(u <= p.cumprob[i])) ++i;         // :356-360
    if (it > 1ull && gn == i)
      for (int j = tid; j < dp; j += 256) gh[j] = fresh[j];
  }
  if (p.split_server) return;                  // server stage runs in shared_server_kernel
  // ------------------------------------------------------------------ server (last CTA)
  __threadfence();
  __syncthreads();
  if (tid == 0) ticket = atomicAdd(p.counters, 1u);
  __syncthreads();
  if (ticket != (unsigned)(p.N - 1)) return;
  __threadfence();
  double f = 0.0;
  for (int i = tid; i < p.N; i += 256) f += __ldcg(p.fw + i);
  f = block_sum(f, red);
  double dn = 0.0;
  for (int j = tid; j < dp; j += 256) {
    double S = 0.0;
    if (p.kind == GADMM_FO_GD) {
      if (it > 1ull) { for (int i = 0; i < p.N; ++i) S += __ldcg(p.Y + (size_t)i * dp + j); }
      else S = p.logistic ? __ldcg(p.Y + j) : (j < p.d ? 1.0 : 0.0);   // ones / worker 1 only
    } else {
      for (int i = 0; i < p.N; ++i) S += __ldcg(p.ghat + (size_t)i * dp + j);
    }
    const double step = p.step * S;
    p.theta[j] -= step;
    dn = fma(step, step, dn);
  }
  dn = block_sum(dn, red);
  if (tid == 0) {
    p.dnorm[it % kSlots] = dn;
    const double loss = fabs(f - p.obj0);
    p.obj_hist[it] = f;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    if (p.kind == GADMM_FO_LAG_PS || p.kind == GADMM_FO_LAG_WK) {
      p.ctrl[3] += (unsigned long long)p.counters[1];
      if (p.counters[1] == 0u && it % 1000ull == 0ull) p.ctrl[3] += 1ull;   // idle bump (:246-251)
      p.comm_hist[it] = (double)p.ctrl[3];
    }
    p.counters[0] = 0u;
    p.counters[1] = 0u;
    if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it;
    p.ctrl[0] = it + 1ull;
  }
}

// ------------------------------------------------- shared-iterate server, parallel + multi-GPU
// grid = dp/64 CTAs x 256 threads (4 worker groups x 64 columns).  Phase 1: column sums of the
// local workers' (stale) gradients; with several GPUs every CTA pushes its 64 partial sums into
// all peers' `xbuf` (NVLink P2P stores), the last CTA adds this GPU's objective partial and upload
// count and raises the per-rank flag on every GPU.  Phase 2: all CTAs wait for all ranks' flags,
// add the partials in rank order (identical result everywhere -> the server step is replicated,
// no broadcast), step theta, and the last CTA closes the iteration.  This is the fused
// reduce + update of SURVEY.md K7.
__global__ void __launch_bounds__(256)
shared_server_kernel(const GadmmShared p) {
  __shared__ double part[4][64];
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int tid = threadIdx.x, col = tid & 63, grp = tid >> 6, dp = p.dp;
  const int j = blockIdx.x * 64 + col;
  const size_t row = (size_t)dp + 64;
  const size_t slot = (size_t)(it & 1ull) * p.nranks * row;
  // ---- phase 1: local column sums (fixed partition of the workers into 4 contiguous groups)
  double S = 0.0;
  const int n0 = (p.N * grp) / 4, n1 = (p.N * (grp + 1)) / 4;
  if (p.kind == GADMM_FO_GD) {
    if (it > 1ull) { for (int n = n0; n < n1; ++n) S += __ldcg(p.Y + (size_t)n * dp + j); }
    else if (grp == 0 && p.gid0 == 0) S = p.logistic ? __ldcg(p.Y + j) : (j < p.d ? 1.0 : 0.0);
  } else {
    for (int n = n0; n < n1; ++n) S += __ldcg(p.ghat + (size_t)n * dp + j);
  }
  part[grp][col] = S;
  __syncthreads();
  S = (part[0][col] + part[1][col]) + (part[2][col] + part[3][col]);
  if (p.nranks > 1) {
    if (grp == 0)
      for (int q = 0; q < p.nranks; ++q) p.xbuf_peer[q][slot + (size_t)p.rank * row + j] = S;
    __threadfence_system();
    __syncthreads();
    if (tid == 0) ticket = atomicAdd(p.counters2, 1u);
    __syncthreads();
    if (ticket == gridDim.x - 1) {              // this GPU's partials are all written
      double f = 0.0;
      for (int i = tid; i < p.N; i += 256) f += __ldcg(p.fw + i);
      f = block_sum(f, red);
      if (tid == 0) {
        p.counters2[0] = 0u;
        for (int q = 0; q < p.nranks; ++q) {
          double* dst = p.xbuf_peer[q] + slot + (size_t)p.rank * row + dp;
          dst[0] = f;
          dst[1] = (double)p.counters[1];
        }
        __threadfence_system();
        for (int q = 0; q < p.nranks; ++q)
          st_release_sys(p.xflag_peer[q] + (it & 1ull) * p.nranks + p.rank, it);
      }
    }
    // ---- phase 2: wait for every rank, add in rank order
    if (tid == 0)
      for (int q = 0; q < p.nranks; ++q) wait_flag_ge(p.xflag + (it & 1ull) * p.nranks + q, it);
    __syncthreads();
    S = 0.0;
    for (int q = 0; q < p.nranks; ++q) S += __ldcg(p.xbuf + slot + (size_t)q * row + j);
  }
  double dn = 0.0;
  if (grp == 0) {
    const double step = p.step * S;
    p.theta[j] -= step;
    dn = step * step;
  }
  dn = block_sum(dn, red);
  if (tid == 0) {
    p.dnpart[blockIdx.x] = dn;
    __threadfence();
    ticket = atomicAdd(p.counters2 + 1, 1u);
  }
  __syncthreads();
  if (ticket != gridDim.x - 1) return;
  // ---- last CTA: close the iteration
  double f = 0.0;
  if (p.nranks == 1)
    for (int i = tid; i < p.N; i += 256) f += __ldcg(p.fw + i);
  f = block_sum(f, red);
  if (tid == 0) {
    double dtot = 0.0;
    for (int b2 = 0; b2 < (int)gridDim.x; ++b2) dtot += __ldcg(p.dnpart + b2);
    unsigned long long up = p.counters[1];
    if (p.nranks > 1) {
      f = 0.0;
      up = 0ull;
      for (int q = 0; q < p.nranks; ++q) {
        f += __ldcg(p.xbuf + slot + (size_t)q * row + dp);
        up += (unsigned long long)__ldcg(p.xbuf + slot + (size_t)q * row + dp + 1);
      }
    }
    p.dnorm[it % kSlots] = dtot;
    const double loss = fabs(f - p.obj0);
    p.obj_hist[it] = f;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    if (p.kind == GADMM_FO_LAG_PS || p.kind == GADMM_FO_LAG_WK) {
      p.ctrl[3] += up;
      if (up == 0ull && it % 1000ull == 0ull) p.ctrl[3] += 1ull;            // idle bump
      p.comm_hist[it] = (double)p.ctrl[3];
    }
    p.counters[0] = 0u;
    p.counters[1] = 0u;
    p.counters2[1] = 0u;
    if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it;
    p.ctrl[0] = it + 1ull;
  }
}

// ------------------------------------------------------ chain halo + objective exchange (K8)
// grid = 3 CTAs.  CTA 0: this GPU's objective partial -> every GPU's fx ring (P2P stores + release
// flag), wait for all ranks, total -> ftot[it & 1] (same order everywhere -> same stop decision).
// CTA 1 / CTA 2 (push_y): first / last local gradient row -> the left / right neighbour GPU's Ytab
// (after the neighbour has posted its partial of this iteration, i.e. has consumed the old row).
__global__ void __launch_bounds__(256)
own_exchange_kernel(const GadmmOwn p, const int push_y) {
  __shared__ double red[32];
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  const int tid = threadIdx.x, dp = p.dp;
  if (blockIdx.x == 0) {
    double f = 0.0;
    for (int i = tid; i < p.N; i += 256) f += __ldcg(p.fw + i);
    f = block_sum(f, red);
    if (tid == 0) {
      const size_t ring = (size_t)(it & 1ull) * p.nranks;
      for (int q = 0; q < p.nranks; ++q) p.fx_peer[q][ring + p.rank] = f;
      __threadfence_system();
      for (int q = 0; q < p.nranks; ++q) st_release_sys(p.fxflag_peer[q] + ring + p.rank, it);
      double tot = 0.0;
      for (int q = 0; q < p.nranks; ++q) {
        wait_flag_ge(p.fxflag + ring + q, it);
        tot += *reinterpret_cast<volatile double*>(p.fx + ring + q);
      }
      p.ftot[it & 1ull] = tot;
    }
    return;
  }
  if (!push_y) return;
  const bool left = blockIdx.x == 1;
  const int gsrc = left ? p.gid0 : p.gid0 + p.N - 1;     // global row to publish
  const int peer = left ? p.rank - 1 : p.rank + 1;
  if (peer < 0 || peer >= p.nranks) return;
  // Back-pressure: the peer's halo row is single-buffered, and its dgd_update(it-1) reads it.
  // The peer posts its objective partial of iteration `it` (CTA 0 of this kernel over there) only
  // after that update has completed in its stream, so waiting for the post makes the overwrite
  // safe.  No cycle: the post does not depend on anything this GPU does in iteration `it`.
  if (tid == 0) wait_flag_ge(p.fxflag + (size_t)(it & 1ull) * p.nranks + peer, it);
  __syncthreads();
  const double* src = p.Ytab + (size_t)gsrc * dp;
  double* dst = p.Ytab_peer[peer] + (size_t)gsrc * dp;
  for (int j = tid; j < dp; j += 256) dst[j] = src[j];
  __threadfence_system();
  __syncthreads();
  if (tid == 0) st_release_sys(p.rowflag_peer[peer] + gsrc, it);
}

// -------------------------------------------------------------------------- DGD update (K8)
// theta_n -= step * mean(g over {n, n+-1})  (1/2 at the ends, 1/3 inside), g = ones at
// iteration 1 (GD_DGD_LAG.m:129-180).  The objective is the one *before* the update (fw).
__global__ void __launch_bounds__(256)
dgd_update_kernel(const GadmmOwn p) {
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int n = blockIdx.x, tid = threadIdx.x, dp = p.dp;
  const int gn = p.gid0 + n, NG = p.Nglob;                 // global position on the chain
  if (p.nranks > 1 && it > 1ull && tid == 0) {             // halo rows pushed by the neighbour GPUs
    if (n == 0 && gn > 0) wait_flag_ge(p.rowflag + gn - 1, it);
    if (n == p.N - 1 && gn < NG - 1) wait_flag_ge(p.rowflag + gn + 1, it);
  }
  __syncthreads();
  const double* Yg = p.Ytab + (size_t)gn * dp;             // == p.Y + n*dp
  for (int j = tid; j < p.d; j += 256) {
    double avg = 1.0;
    if (it > 1ull) {
      const double g = Yg[j];
      if (NG == 1) avg = g;
      else if (gn == 0) avg = 0.5 * (g + __ldcg(Yg + dp + j));
      else if (gn == NG - 1) avg = 0.5 * (g + __ldcg(Yg - dp + j));
      else avg = (g + __ldcg(Yg + dp + j) + __ldcg(Yg - dp + j)) / 3.0;
    }
    p.Theta[(size_t)n * dp + j] -= p.step * avg;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) ticket = atomicAdd(p.counters, 1u);
  __syncthreads();
  if (ticket != (unsigned)(p.N - 1)) return;
  double f = 0.0;
  for (int i = tid; i < p.N; i += 256) f += __ldcg(p.fw + i);
  f = block_sum(f, red);
  if (tid == 0) {
    if (p.nranks > 1) f = p.ftot[it & 1ull];
    const double loss = fabs(f - p.obj0);
    p.obj_hist[it] = f;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    p.counters[0] = 0u;
    if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it;
    p.ctrl[0] = it + 1ull;
  }
}

// ---------------------------------------------------------------- dual averaging update (K8)
// ctrl[0] = index of the sweep about to run (1-based).  fw holds f(Theta) of the iterate produced
// by sweep ctrl[0]-1; it is recorded first (that closes iteration ctrl[0]-1), then -- unless the
// stop test fired -- one sweep along the chain runs: thread j walks n = 0..N-1 with
//   Z_n = [Z_{n+1}(old) | Z_{n-1} | mean of both] + g_n,  theta_n = -step*Z_n
// `jacobi`=0 reproduces the reference's Gauss-Seidel order (Z_prev overwritten in the loop,
// dual_averaging.m:44: worker n sees the NEW Z of n-1); `jacobi`=1 uses the old one.
__global__ void __launch_bounds__(256)
dualavg_update_kernel(const GadmmOwn p, const int update) {
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  __shared__ int s_stop;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  const int tid = threadIdx.x, dp = p.dp, N = p.N, NG = p.Nglob, g0 = p.gid0;
  double f = 0.0;
  for (int i = tid; i < N; i += 256) f += __ldcg(p.fw + i);
  f = block_sum(f, red);                       // identical in every CTA (same order)
  if (p.nranks > 1) f = p.ftot[it & 1ull];     // all-rank objective from own_exchange_kernel
  const double loss = fabs(f - p.obj0);
  if (tid == 0) s_stop = (it > 1ull && loss < p.acc && p.ctrl[2] != 0ull) ? 1 : 0;
  __syncthreads();
  const bool run = update && !s_stop && it <= (unsigned long long)p.max_iter;
  // halo rows of Z: left neighbour's last worker (new value in Gauss-Seidel order, old in Jacobi),
  // right neighbour's first worker (always the previous iterate)
  const bool haloL = p.nranks > 1 && g0 > 0, haloR = p.nranks > 1 && g0 + N < NG;
  const unsigned long long vL = p.jacobi ? it - 1ull : it, vR = it - 1ull;
  if (run && tid == 0) {
    if (haloL) wait_flag_ge(p.rowflag + NG + g0 - 1, vL);
    if (haloR) wait_flag_ge(p.rowflag + NG + g0 + N, vR);
  }
  __syncthreads();
  if (run) {
    const int j = blockIdx.x * 256 + tid;
    if (j < p.d) {
      const double* ZL = p.Ztab + ((size_t)(vL & 1ull) * NG + (g0 - 1)) * dp;
      const double* ZR = p.Ztab + ((size_t)(vR & 1ull) * NG + (g0 + N)) * dp;
      double left_new = haloL ? __ldcg(ZL + j) : 0.0, left_old = left_new;
      double z_first = 0.0, z_last = 0.0;
      for (int n = 0; n < N; ++n) {
        const size_t o = (size_t)n * dp + j;
        const int gn = g0 + n;
        const double zo = p.Z[o];
        const double lft = p.jacobi ? left_old : left_new;
        const double rgt = (n == N - 1) ? (haloR ? __ldcg(ZR + j) : 0.0) : p.Z[o + dp];
        double z;
        if (NG == 1) z = 0.0;
        else if (gn == 0) z = rgt;
        else if (gn == NG - 1) z = lft;
        else z = 0.5 * rgt + 0.5 * lft;
        z += p.Y[o];
        p.Theta[o] = -p.step * z;
        p.Z[o] = z;
        left_new = z;
        left_old = zo;
        if (n == 0) z_first = z;
        z_last = z;
      }
      // publish the boundary rows (version it) into the neighbour GPUs' halo tables
      const size_t so = (size_t)(it & 1ull) * NG;
      if (haloL) p.Ztab_peer[p.rank - 1][(so + g0) * dp + j] = z_first;
      if (haloR) p.Ztab_peer[p.rank + 1][(so + g0 + N - 1) * dp + j] = z_last;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid == 0) ticket = atomicAdd(p.counters, 1u);
  __syncthreads();
  if (ticket != gridDim.x - 1) return;
  if (tid == 0) {
    p.counters[0] = 0u;
    if (run) {
      __threadfence_system();
      if (haloL) st_release_sys(p.rowflag_peer[p.rank - 1] + NG + g0, it);
      if (haloR) st_release_sys(p.rowflag_peer[p.rank + 1] + NG + g0 + N - 1, it);
    }
    if (it > 1ull) {                            // close iteration it-1
      p.obj_hist[it - 1] = f;
      p.loss_hist[it - 1] = loss;
      p.t_hist[it - 1] = globaltimer_ns();
      if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it - 1ull;
    }
    if (run) p.ctrl[0] = it + 1ull;
  }
}

template <typename T>
static int launch_symv_grad(const void* const* Sptr, const double* x, long long xs,
                            const double* b, const double* hyy, double* Y, double* fw, int N,
                            int dp, double* scratch, unsigned int* counters, cudaStream_t st) {
  const int smem = (int)sizeof(SymvSmem<T>);
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (!configured) {
    GADMM_CUDA_OK(cudaFuncSetAttribute(symv_grad_kernel<T>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = 1;
  }
  const int nT = dp / kT, npairs = (nT + 1) / 2;
  symv_grad_kernel<T><<<N * npairs, kSymvThreads, smem, st>>>(Sptr, x, xs, b, hyy, Y, fw, dp,
                                                              scratch, counters);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

int gadmm_symv_grad(const void* const* Sptr, const double* x, long long x_stride, const double* b,
                    const double* hyy, double* Y, double* fw, int N, int dp, int store,
                    double* scratch, unsigned int* counters, void* stream) {
  GADMM_REQUIRE(dp > 0 && dp % kT == 0, "symv_grad: dp=%d must be a positive multiple of 64", dp);
  if (N <= 0) return 0;
  if (store == GADMM_STORE_F64)
    return launch_symv_grad<double>(Sptr, x, x_stride, b, hyy, Y, fw, N, dp, scratch, counters,
                                    (cudaStream_t)stream);
  if (store == GADMM_STORE_F32)
    return launch_symv_grad<float>(Sptr, x, x_stride, b, hyy, Y, fw, N, dp, scratch, counters,
                                   (cudaStream_t)stream);
  return set_error(-1, "symv_grad: bad store %d", store);
}

int gadmm_logistic_grad(const double* X, const double* y, const double* th, long long th_stride,
                        double lam, int N, int s, int d, int dp, double* Y, double* fw,
                        double* wout, double* wbuf, void* stream) {
  GADMM_REQUIRE(N > 0 && s > 0 && d > 0 && dp >= d, "logistic_grad: bad shape");
  logistic_grad_kernel<<<N, 256, 0, (cudaStream_t)stream>>>(X, y, th, th_stride, lam, s, d, dp, Y,
                                                            fw, wout, wbuf);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_shared_update(const GadmmShared* p, void* stream) {
  GADMM_REQUIRE(p && p->N > 0, "shared_update: N must be positive");
  GADMM_REQUIRE(p->split_server || p->nranks == 1, "shared_update: multi-GPU needs split_server");
  shared_update_kernel<<<p->N, 256, 0, (cudaStream_t)stream>>>(*p);
  if (p->split_server)
    shared_server_kernel<<<p->dp / 64, 256, 0, (cudaStream_t)stream>>>(*p);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_own_exchange(const GadmmOwn* p, int push_y, void* stream) {
  GADMM_REQUIRE(p && p->N > 0, "own_exchange: N must be positive");
  if (p->nranks <= 1) return 0;
  own_exchange_kernel<<<3, 256, 0, (cudaStream_t)stream>>>(*p, push_y);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_dgd_update(const GadmmOwn* p, void* stream) {
  GADMM_REQUIRE(p && p->N > 0, "dgd_update: N must be positive");
  dgd_update_kernel<<<p->N, 256, 0, (cudaStream_t)stream>>>(*p);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_dualavg_update(const GadmmOwn* p, int update, void* stream) {
  GADMM_REQUIRE(p && p->N > 0, "dualavg_update: N must be positive");
  dualavg_update_kernel<<<(p->d + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*p, update);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
