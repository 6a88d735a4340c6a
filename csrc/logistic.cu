// Logistic-regression GADMM local solves (K9 of SURVEY.md section 2.4).
//
// logistic_half_gd_kernel: the reference's inexact local solve -- at most 100 fixed-step gradient
// steps on  lam/2||x||^2 + sum log(1+exp(-y.*(X x))) - l_{n-1}'x + l_n'x + penalties, with the
// penalty gradient terms rho(theta_n - theta_{n+-1}) evaluated once at entry and frozen
// (group_ADMM_logistic_GD.m:24-52,75-90), warm start from theta_n, exit when every
// |dx_j| < 1e-4 (logReg_GD.m:3-25).  One CTA per worker runs the whole inner loop on chip:
// x, the frozen linear term and the per-sample weights live in shared memory, X_n streams from
// L2.  The exact (Newton) solve used for the scaled configs is orchestrated in
// gadmm_b200/engine/logistic.py from the Gram / Cholesky / SYMV kernels.
//
// logistic_dual_obj_kernel: lambda_n += rho(theta_n - theta_{n+1}) (:99-103), objective sum of
// the per-worker losses produced by gadmm_logistic_grad (:109-115), stop test, iteration++.
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

__global__ void __launch_bounds__(256)
logistic_half_gd_kernel(const GadmmLogistic p, const int color) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double* x = reinterpret_cast<double*>(dsm);        // [d]
  double* cst = x + p.dp;                            // [d]  -C1 + C2 + t1 + t2
  double* w = cst + p.dp;                            // [s]
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int n = p.n0 + blockIdx.x;                   // global worker; shards are indexed locally
  if ((n & 1) != color) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, d = p.d, s = p.s, dp = p.dp;
  const double* Xn = p.X + (size_t)blockIdx.x * s * d;
  const double* yn = p.y + (size_t)blockIdx.x * s;
  double* th = p.theta + (size_t)n * dp;
  const bool hasL = n > 0, hasR = n < p.N - 1;
  for (int j = tid; j < d; j += 256) {
    const double t = th[j];
    double c = 0.0;
    if (hasL) c += -p.lam[(size_t)(n - 1) * dp + j] + p.rho * (t - p.theta[(size_t)(n - 1) * dp + j]);
    if (hasR) c += p.lam[(size_t)n * dp + j] + p.rho * (t - p.theta[(size_t)(n + 1) * dp + j]);
    x[j] = t;
    cst[j] = c;
  }
  __syncthreads();
  int steps = 0;
  for (int k = 0; k < p.max_inner; ++k) {
    for (int r = warp; r < s; r += 8) {
      double z = 0.0;
      for (int j = lane; j < d; j += 32) z = fma(Xn[(size_t)r * d + j], x[j], z);
      z = warp_sum(z);
      if (lane == 0) w[r] = yn[r] / (1.0 + exp(yn[r] * z));      // logReg_GD.m:9
    }
    __syncthreads();
    int small = 1;
    for (int j = tid; j < d; j += 256) {
      double g = 0.0;
      for (int r = 0; r < s; ++r) g = fma(Xn[(size_t)r * d + j], w[r], g);
      g = -g + p.lam_log * x[j] + cst[j];
      const double dx = p.gd_step * g;
      if (!(fabs(dx) < p.inner_tol)) small = 0;
      x[j] -= dx;                        // only this thread touches x[j] in this phase
    }
    ++steps;
    if (__syncthreads_and(small)) break;                      // all(abs(dx) < 1e-4), :21-23
  }
  for (int j = tid; j < d; j += 256) th[j] = x[j];
  if (tid == 0 && p.inner_steps) atomicAdd(p.inner_steps, (unsigned long long)steps);
}

__global__ void __launch_bounds__(256)
logistic_dual_obj_kernel(const GadmmLogistic p) {
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int n = blockIdx.x, tid = threadIdx.x, dp = p.dp;
  if (n < p.N - 1)
    for (int j = tid; j < p.d; j += 256)
      p.lam[(size_t)n * dp + j] += p.rho * (p.theta[(size_t)n * dp + j] - p.theta[(size_t)(n + 1) * dp + j]);
  __threadfence();
  __syncthreads();
  if (tid == 0) ticket = atomicAdd(p.counters, 1u);
  __syncthreads();
  if (ticket != (unsigned)(p.N - 1)) return;
  double f = 0.0;
  for (int i = tid; i < p.N; i += 256) f += __ldcg(p.fw + i);
  f = block_sum(f, red);
  if (tid == 0) {
    const double loss = fabs(f - p.obj0);
    p.obj_hist[it] = f;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    p.counters[0] = 0u;
    if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it;
    p.ctrl[0] = it + 1ull;
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

int gadmm_logistic_half_gd(const GadmmLogistic* p, int color, void* stream) {
  GADMM_REQUIRE(p && p->N > 0 && p->d > 0 && p->s > 0, "logistic_half_gd: bad shape");
  const int smem = (2 * p->dp + p->s) * (int)sizeof(double);
  GADMM_REQUIRE(smem <= 200 * 1024, "logistic_half_gd: d=%d, s=%d needs %d B of shared memory "
                "(> 200 KiB); use the Newton path", p->d, p->s, smem);
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (smem > configured) {
    GADMM_CUDA_OK(cudaFuncSetAttribute(logistic_half_gd_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = smem;
  }
  GadmmLogistic q = *p;
  if (q.nloc <= 0) { q.n0 = 0; q.nloc = q.N; }
  GADMM_REQUIRE(q.n0 >= 0 && q.n0 + q.nloc <= q.N, "logistic_half_gd: bad worker range");
  logistic_half_gd_kernel<<<q.nloc, 256, smem, (cudaStream_t)stream>>>(q, color);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_logistic_dual_obj(const GadmmLogistic* p, void* stream) {
  GADMM_REQUIRE(p && p->N > 0, "logistic_dual_obj: N must be positive");
  logistic_dual_obj_kernel<<<p->N, 256, 0, (cudaStream_t)stream>>>(*p);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
