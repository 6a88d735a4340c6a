// Persistent chain-GADMM kernel for small problems (reference shapes: N = 24..50 workers, d = 50).
//
// At the reference's own sizes one iteration of the general path is three launch-bound kernels
// (~30 us).  Here the whole solve is ONE cooperative launch: one CTA per worker, co-resident on
// the GPU for all iterations.  (G_n + c rho I)^-1, b_n, the node dual and the worker's own theta
// stay in shared memory; the only global traffic per half-iteration is the neighbours' theta
// (2 x d doubles) and a version flag.  Head/tail ordering needs no grid barrier: a CTA waits only
// for its two neighbours' flags (ld.acquire.gpu), exactly the dependency structure of
// group_ADMM_closedForm.m:13-57 / :61-92.  The objective (:96-101) is summed in a fixed order by
// whichever CTA arrives last for that iteration; a grid-wide barrier every kSyncEvery iterations
// lets all CTAs agree on the stop decision (:103-108), so the state stays consistent.
// "persistent cooperative kernels ... flags are the only
// synchronisation".
#include "../include/gadmm_b200.h"
#include "common.cuh"

#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace gadmm {

constexpr int kSyncEvery = 8;
constexpr int kObjRing = 32;

__global__ void __launch_bounds__(256, 1)
gadmm_persistent_kernel(const GadmmChain c, const int n_iter, double* __restrict__ objring,
                        unsigned int* __restrict__ ringcnt) {
  extern __shared__ __align__(16) unsigned char dsm[];
  const int dp = c.dp, tid = threadIdx.x, k = blockIdx.x, nT = dp / 64;
  double* M = reinterpret_cast<double*>(dsm);       // [dp][dp] full symmetric
  double* b = M + (size_t)dp * dp;                  // [dp]
  double* nu = b + dp;
  double* th = nu + dp;                             // own theta
  double* r = th + dp;
  double* part = r + dp;                            // [4][dp] matvec partials
  double* hist = part + 4 * dp;                     // [kSyncEvery][2][dp] (theta, nu) of the last
                                                    // iterations: exact roll-back to the stop iteration
  __shared__ double red[32];
  __shared__ unsigned int s_ticket;
  cg::grid_group grid = cg::this_grid();

  const int g = c.gid[k], l = c.left[k], rr = c.right[k], color = c.color[k];
  const double rho = c.rho, crho = c.rho * (double)c.cmul[k];
  // ---- stage the worker's state in shared memory (unpack the tile-packed inverse)
  const double* P = reinterpret_cast<const double*>(c.Mptr[k]);
  for (int idx = tid; idx < dp * dp; idx += 256) {
    const int i = idx / dp, j = idx % dp;
    const int hi = i >= j ? i : j, lo = i >= j ? j : i;            // lower triangle is stored
    const int I = hi / 64, J = lo / 64;
    const int ti = (I == J) ? (i % 64) : (hi % 64), tj = (I == J) ? (j % 64) : (lo % 64);
    M[idx] = P[((size_t)I * (I + 1) / 2 + J) * 4096 + ti * 64 + tj];
  }
  unsigned long long it = c.ctrl[0];
  const size_t slot0 = (size_t)((it - 1ull) & 1ull) * c.Nglob;
  for (int j = tid; j < dp; j += 256) {
    b[j] = c.b[(size_t)k * dp + j];
    nu[j] = c.nu[(size_t)k * dp + j];
    th[j] = c.theta[(slot0 + g) * dp + j];
  }
  __syncthreads();
  const unsigned long long it_end = it + (unsigned long long)n_iter;
  const int nparts = 256 / dp > 0 ? (256 / dp > 4 ? 4 : 256 / dp) : 1;     // threads per row group
  bool stopped = (c.ctrl[2] != 0ull && c.ctrl[1] != 0ull);

  for (; it < it_end && it <= (unsigned long long)c.max_iter && !stopped; ++it) {
    // ------------------------------------------------------------------ solve (head or tail)
    const unsigned long long need = (color == 0) ? it - 1ull : it;
    const size_t sin = (size_t)(need & 1ull) * c.Nglob, sout = (size_t)(it & 1ull) * c.Nglob;
    if (tid == 0) {
      if (l >= 0) while (ld_acquire_gpu(c.flags + l) < need) {}
      if (rr >= 0) while (ld_acquire_gpu(c.flags + rr) < need) {}
    }
    __syncthreads();
    for (int j = tid; j < dp; j += 256) {
      double t = 0.0;
      if (l >= 0) t += __ldcg(c.theta + (sin + l) * dp + j);
      if (rr >= 0) t += __ldcg(c.theta + (sin + rr) * dp + j);
      r[j] = fma(rho, t, b[j] - nu[j]);
    }
    __syncthreads();
    // theta = M r : thread (q, i) accumulates rows j = q, q+nparts, ... of column i (M symmetric
    // -> conflict-free reads of M[j][i]); partials are combined in a fixed order
    if (tid < nparts * dp) {
      const int i = tid % dp, q = tid / dp;
      double acc = 0.0;
      for (int j = q; j < dp; j += nparts) acc = fma(M[(size_t)j * dp + i], r[j], acc);
      part[q * dp + i] = acc;
    }
    __syncthreads();
    double s1 = 0.0, s2 = 0.0;
    for (int i = tid; i < dp; i += 256) {
      double v = part[i];
      for (int q = 1; q < nparts; ++q) v += part[q * dp + i];
      th[i] = v;
      c.theta[(sout + g) * dp + i] = v;
      s1 = fma(v, r[i] - crho * v, s1);
      s2 = fma(b[i], v, s2);
    }
    s1 = block_sum(s1, red);
    s2 = block_sum(s2, red);
    if (tid == 0) {
      __threadfence();
      st_release_gpu(c.flags + g, it);
    }
    // ------------------------------------------------------------------ dual update
    if (tid == 0) {
      if (l >= 0) while (ld_acquire_gpu(c.flags + l) < it) {}
      if (rr >= 0) while (ld_acquire_gpu(c.flags + rr) < it) {}
    }
    __syncthreads();
    for (int j = tid; j < dp; j += 256) {
      double upd = 0.0;
      if (rr >= 0) upd = upd + rho * (th[j] - __ldcg(c.theta + (sout + rr) * dp + j));
      if (l >= 0) upd = upd - rho * (__ldcg(c.theta + (sout + l) * dp + j) - th[j]);
      nu[j] += upd;
      double* h = hist + (size_t)(it % kSyncEvery) * 2 * dp;
      h[j] = th[j];
      h[dp + j] = nu[j];
    }
    // ------------------------------------------------------------------ objective (last arriver)
    const int ring = (int)(it % kObjRing);
    if (tid == 0) {
      objring[(size_t)ring * c.Nloc + k] = 0.5 * s1 - s2 + c.hyy[k];
      __threadfence();
      s_ticket = atomicAdd(ringcnt + ring, 1u);
    }
    __syncthreads();
    if (s_ticket == (unsigned)(c.Nloc - 1)) {
      __threadfence();
      double f = 0.0;
      for (int i = tid; i < c.Nloc; i += 256) f += __ldcg(objring + (size_t)ring * c.Nloc + i);
      f = block_sum(f, red);
      if (tid == 0) {
        ringcnt[ring] = 0u;
        const double loss = fabs(f - c.obj0);
        c.obj_hist[it] = f;
        c.loss_hist[it] = loss;
        c.t_hist[it] = globaltimer_ns();
        if (loss < c.acc) atomicMin(c.ctrl + 3, it);        // first crossing (ctrl[3] = +inf initially)
      }
    }
    // ------------------------------------------------------------------ agree on stopping
    if ((it % kSyncEvery) == 0ull || it + 1ull == it_end || it == (unsigned long long)c.max_iter) {
      __threadfence();
      grid.sync();
      const unsigned long long hit = *reinterpret_cast<volatile unsigned long long*>(c.ctrl + 3);
      if (c.ctrl[2] != 0ull && hit <= it) stopped = true;
    }
  }
  // ---- write the state back so that the general path / checkpointing can continue from here.
  // The stop is agreed on only every kSyncEvery iterations, so CTAs may have run up to
  // kSyncEvery-1 iterations past the first crossing `hit`: roll back to the state of iteration
  // `hit` (the reference breaks there, group_ADMM_closedForm.m:105-108), so that state() / resume
  // and the general path see exactly what they would have produced.
  __syncthreads();
  const unsigned long long hit = *reinterpret_cast<volatile unsigned long long*>(c.ctrl + 3);
  const bool rolled = stopped && c.ctrl[2] != 0ull && hit != ~0ull && hit + 1ull < it;
  if (rolled) {
    const double* h = hist + (size_t)(hit % kSyncEvery) * 2 * dp;
    const size_t so = (size_t)(hit & 1ull) * c.Nglob;
    for (int j = tid; j < dp; j += 256) {
      nu[j] = h[dp + j];
      c.theta[(so + g) * dp + j] = h[j];
    }
    if (tid == 0) c.flags[g] = hit;
  }
  __syncthreads();
  for (int j = tid; j < dp; j += 256) c.nu[(size_t)k * dp + j] = nu[j];
  if (k == 0 && tid == 0) {
    if (c.ctrl[2] != 0ull && hit != ~0ull && hit < it && c.ctrl[1] == 0ull) c.ctrl[1] = hit;
    c.ctrl[0] = rolled ? hit + 1ull : it;
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

/* Returns 0 when the persistent path can run this problem (single GPU, fp64 tiles, dp <= 128,
 * one CTA per worker co-resident), non-zero otherwise. */
int gadmm_chain_persistent_supported(const GadmmChain* c) {
  if (!c || c->nranks != 1 || c->store != GADMM_STORE_F64 || c->dp > 128) return 1;
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 1;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 1;
  return c->Nloc <= sms ? 0 : 1;
}

/* objring: kObjRing*Nloc doubles, ringcnt: kObjRing uint32 (zero-initialised); ctrl[3] must hold
 * ~0ull before the first call (first-crossing register). */
int gadmm_chain_persistent(const GadmmChain* c, int n_iter, double* objring, unsigned int* ringcnt,
                           void* stream) {
  GADMM_REQUIRE(gadmm_chain_persistent_supported(c) == 0, "persistent path: unsupported problem");
  const int dp = c->dp;
  const size_t smem = ((size_t)dp * dp + 4 * (size_t)dp + 4 * (size_t)dp +
                       2 * (size_t)kSyncEvery * dp) * sizeof(double);
  GADMM_CUDA_OK(cudaFuncSetAttribute(gadmm_persistent_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  GadmmChain cc = *c;
  void* args[] = {(void*)&cc, (void*)&n_iter, (void*)&objring, (void*)&ringcnt};
  GADMM_CUDA_OK(cudaLaunchCooperativeKernel((void*)gadmm_persistent_kernel, dim3(c->Nloc), dim3(256),
                                            args, smem, (cudaStream_t)stream));
  return 0;
}

}  // extern "C"
