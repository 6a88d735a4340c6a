// Standard (star / parameter-server) ADMM, closed form -- K6 of SURVEY.md section 2.4.
//
//   workers n < N:  theta_n <- (G_n + rho I)^-1 (b_n - lambda_n + rho z)          standared_ADMM.m:42
//   server  N    :  z <- (G_N + (N-1) rho I)^-1 (b_N + sum lambda_n + rho sum theta_n)   :66-73
//   duals        :  lambda_n += rho (theta_n - z)                                          :84-88
//
// star_workers_kernel: streamed packed-symmetric SYMV per worker; the dual update of the previous
// iteration is folded into the RHS assembly (r = b - [lambda + rho(theta - z)] + rho z) and made
// permanent by the worker's finisher CTA.
// star_reduce_kernel: u = sum_n (lambda_n + rho theta_n), one CTA per 64 columns, fixed order,
// pushed -- together with this GPU's objective partial -- into every GPU's `ubuf` with NVLink P2P
// stores + st.release.sys.
// star_server_kernel + star_zfinish_kernel: replicated on every GPU (no broadcast of z): wait for
// all ranks' partials with ld.acquire.sys, stream the server inverse with npairs*nsplit CTAs,
// finish z per 64 entries, add the server shard's objective, run the stop test.
#include "../include/gadmm_b200.h"
#include "symv_core.cuh"

namespace gadmm {

__device__ __forceinline__ double mm_ld_reduce_add(const double* mc) {
  double v;   // in-switch (NVLS) sum of the same offset on every GPU; SASS: LDGMC.E.ADD.F64
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(v) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st(double* mc, double v) {   // broadcast store to every GPU
  asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(mc), "d"(v) : "memory");
}

constexpr int kUPad = 64;   // ubuf row = dp + kUPad doubles; [dp] carries the objective partial

struct StarWorkerX {
  const double* b;
  const double* lam;
  const double* th;
  const double* z;
  double rho;
  __device__ __forceinline__ double lam_new(int j) const {
    return lam[j] + rho * (th[j] - z[j]);
  }
  __device__ __forceinline__ double operator()(int j) const {
    return b[j] - lam_new(j) + rho * z[j];
  }
  struct Raw { double b, lam, th, z; };
  __device__ __forceinline__ Raw load(int j) const { return Raw{b[j], lam[j], th[j], z[j]}; }
  __device__ __forceinline__ double fin(const Raw& v) const {
    return v.b - (v.lam + rho * (v.th - v.z)) + rho * v.z;
  }
};

template <typename T>
__global__ void __launch_bounds__(kSymvThreads, 2)
star_workers_kernel(const GadmmStar p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SymvSmem<T>& sm = *reinterpret_cast<SymvSmem<T>*>(smem_raw);
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int dp = p.dp, nT = dp / kT, npairs = (nT + 1) / 2;
  const int k = blockIdx.x / npairs, pr = blockIdx.x % npairs;
  StarWorkerX xf{p.b + (size_t)k * dp, p.lam + (size_t)k * dp, p.theta + (size_t)k * dp, p.z, p.rho};
  double* rowpart = p.rowpart + (size_t)k * dp;
  double* colpart = p.colpart + (size_t)k * colpart_elems(dp);
  const bool last = symv_cta<T>(sm, reinterpret_cast<const T*>(p.Mptr[k]), nT, pr, xf, rowpart,
                                colpart, p.counters + k, npairs);
  if (!last) return;
  // ---- finisher of worker k
  double* lam = p.lam + (size_t)k * dp;
  double* th = p.theta + (size_t)k * dp;
  double* U = p.U + (size_t)k * dp;
  double s1 = 0.0, s2 = 0.0;
  for (int j = threadIdx.x; j < dp; j += blockDim.x) {
    const double ln = xf.lam_new(j);
    const double r = xf.b[j] - ln + p.rho * p.z[j];
    const double tn = symv_gather(rowpart, colpart, nT, j);
    lam[j] = ln;
    th[j] = tn;
    U[j] = ln + p.rho * tn;
    s1 = fma(tn, r - p.rho * tn, s1);
    s2 = fma(xf.b[j], tn, s2);
  }
  double* red = &sm.colbuf[0][0][0];
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  if (threadIdx.x == 0) p.fw[k] = 0.5 * s1 - s2 + p.hyy[k];
}

// ---- u = sum_n (lambda_n + rho theta_n) over this GPU's workers, one CTA per 64 columns (256
// threads: 4 groups of workers x 64 columns, combined in a fixed order), pushed -- together with
// the objective partial -- into every GPU's `ubuf` with NVLink P2P stores; the last CTA raises
// the flags.  (NVLS mode: the partial stays local, the server sums in the switch.)
__global__ void __launch_bounds__(256)
star_reduce_kernel(const GadmmStar p) {
  __shared__ double acc[4][64];
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int dp = p.dp, tid = threadIdx.x, t = tid & 63, w = tid >> 6, j = blockIdx.x * 64 + t;
  double u = 0.0;
  for (int i = w; i < p.Nw; i += 4) u += __ldcg(p.U + (size_t)i * dp + j);
  acc[w][t] = u;
  __syncthreads();
  const size_t row = (size_t)dp + kUPad;
  if (w == 0) {
    u = (acc[0][t] + acc[1][t]) + (acc[2][t] + acc[3][t]);
    if (p.nvls) {
      p.ubuf[(size_t)(it & 1ull) * p.nranks * row + j] = u;
    } else {
      const size_t slot = ((size_t)(it & 1ull) * p.nranks + p.rank) * row;
      for (int q = 0; q < p.nranks; ++q) p.ubuf_peer[q][slot + j] = u;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid == 0) ticket = atomicAdd(p.counters + p.Nw + 2, 1u);
  __syncthreads();
  if (ticket != gridDim.x - 1) return;
  __threadfence();
  double fpart = 0.0;
  for (int i = tid; i < p.Nw; i += blockDim.x) fpart += __ldcg(p.fw + i);
  fpart = block_sum(fpart, red);
  if (tid == 0) {
    if (p.nvls) {
      p.ubuf[(size_t)(it & 1ull) * p.nranks * row + dp] = fpart;
    } else {
      const size_t slot = ((size_t)(it & 1ull) * p.nranks + p.rank) * row;
      for (int q = 0; q < p.nranks; ++q) p.ubuf_peer[q][slot + dp] = fpart;
    }
    p.counters[p.Nw + 2] = 0u;
    __threadfence_system();
    for (int q = 0; q < p.nranks; ++q)
      st_release_sys(p.uflag_peer[q] + (it & 1ull) * p.nranks + p.rank, it);
  }
}

struct StarServerX {
  const double* bs;
  const double* ubuf;   // this iteration's slot base: [nranks][dp + kUPad]
  const double* mc;     // multicast alias of the slot base (NVLS mode) or null
  int nranks;
  size_t row;
  __device__ __forceinline__ double usum(int j) const {
    if (mc) return mm_ld_reduce_add(mc + j);
    double u = 0.0;
    for (int q = 0; q < nranks; ++q) u += __ldcg(ubuf + q * row + j);   // fixed rank order
    return u;
  }
  __device__ __forceinline__ double operator()(int j) const { return bs[j] + usum(j); }
};

// Server solve, part 1: the packed inverse of the server shard is streamed by npairs*nsplit CTAs
// (a single CTA sustains ~25 GB/s, so the tile columns of every tile-row pair are divided over
// `nsplit` CTAs, as in the chain kernel).  Partial sums go to the server slot of rowpart/colpart.
template <typename T>
__global__ void __launch_bounds__(kSymvThreads, 2)
star_server_kernel(const GadmmStar p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SymvSmem<T>& sm = *reinterpret_cast<SymvSmem<T>*>(smem_raw);
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int dp = p.dp, nT = dp / kT, npairs = (nT + 1) / 2, nsp = p.nsplit > 0 ? p.nsplit : 1;
  if (p.nvls && p.rank != p.server_rank) return;   // NVLS mode: z arrives by multicast (zfinish)
  if (threadIdx.x == kConsumerWarps * 32)
    for (int q = 0; q < p.nranks; ++q) wait_flag_ge(p.uflag + (it & 1ull) * p.nranks + q, it);
  __syncthreads();
  const size_t row = (size_t)dp + kUPad;
  const size_t slot_off = (size_t)(it & 1ull) * p.nranks * row;
  StarServerX xf{p.bs, p.ubuf + slot_off, p.nvls ? p.ubuf_mc + slot_off : nullptr, p.nranks, row};
  const int pair = blockIdx.x / nsp, part = blockIdx.x % nsp;
  double* rowpart = p.rowpart + (size_t)p.Nw * dp;
  double* colpart = p.colpart + (size_t)p.Nw * colpart_elems(dp);
  symv_cta<T>(sm, reinterpret_cast<const T*>(p.Ms), nT, pair, xf, rowpart + (size_t)part * dp,
              colpart, p.counters + p.Nw + 1, npairs * nsp, part, nsp);
}

// Server solve, part 2: one 64-thread CTA per 64 entries of z adds the partial sums in a fixed
// order, writes z (NVLS: multicast store to every GPU), and the last CTA closes the iteration:
// total objective (standared_ADMM.m:95-100), stop test, clock, iteration counter.
__global__ void __launch_bounds__(64)
star_zfinish_kernel(const GadmmStar p) {
  __shared__ double red[4];
  __shared__ unsigned int ticket;
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  if (it > (unsigned long long)p.max_iter) return;
  const int dp = p.dp, nT = dp / kT, tid = threadIdx.x, nsp = p.nsplit > 0 ? p.nsplit : 1;
  const size_t row = (size_t)dp + kUPad;
  const size_t slot_off = (size_t)(it & 1ull) * p.nranks * row;
  const double* ub = p.ubuf + slot_off;
  if (p.nvls && p.rank != p.server_rank) {
    // NVLS mode, non-server GPU: z and the objective arrive by multicast store from the server
    if (blockIdx.x != 0 || tid != 0) return;
    wait_flag_ge(p.zflag, it);
    const double total = *reinterpret_cast<volatile double*>(p.zobj + (it & 3ull));
    const double loss = fabs(total - p.obj0);
    p.obj_hist[it] = total;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it;
    p.ctrl[0] = it + 1ull;
    return;
  }
  StarServerX xf{p.bs, ub, p.nvls ? p.ubuf_mc + slot_off : nullptr, p.nranks, row};
  const double* rowpart = p.rowpart + (size_t)p.Nw * dp;
  const double* colpart = p.colpart + (size_t)p.Nw * colpart_elems(dp);
  const int J = blockIdx.x, j = J * kT + tid;
  const double zn = symv_gather(rowpart, colpart, nT, j, nsp, dp);
  const double r = xf(j);
  if (p.nvls) mm_st(p.z_mc + j, zn); else p.z[j] = zn;      // NVLS: broadcast z to every GPU
  const double crho = p.rho * (double)(p.N - 1);
  double s1 = zn * (r - crho * zn), s2 = p.bs[j] * zn;
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  if ((tid & 31) == 0) { red[(tid >> 5) * 2] = s1; red[(tid >> 5) * 2 + 1] = s2; }
  __syncthreads();
  if (tid == 0) {
    p.zblk[2 * J] = red[0] + red[2];
    p.zblk[2 * J + 1] = red[1] + red[3];
    __threadfence_system();
    ticket = atomicAdd(p.counters + p.Nw + 3, 1u);
  }
  __syncthreads();
  if (ticket != (unsigned)(nT - 1)) return;
  __threadfence();
  double a1 = 0.0, a2 = 0.0;
  for (int q = tid; q < nT; q += 64) {           // fixed partition + fixed tree: deterministic
    a1 += __ldcg(p.zblk + 2 * q);
    a2 += __ldcg(p.zblk + 2 * q + 1);
  }
  a1 = warp_sum(a1);
  a2 = warp_sum(a2);
  __syncthreads();
  if ((tid & 31) == 0) { red[(tid >> 5) * 2] = a1; red[(tid >> 5) * 2 + 1] = a2; }
  __syncthreads();
  if (tid == 0) {
    p.counters[p.Nw + 3] = 0u;
    double total = 0.5 * (red[0] + red[2]) - (red[1] + red[3]) + p.hyy_s;
    if (p.nvls) total += mm_ld_reduce_add(p.ubuf_mc + slot_off + dp);
    else for (int q = 0; q < p.nranks; ++q) total += __ldcg(ub + q * row + dp);
    const double loss = fabs(total - p.obj0);
    p.obj_hist[it] = total;
    p.loss_hist[it] = loss;
    p.t_hist[it] = globaltimer_ns();
    if (p.nvls) {
      mm_st(p.zobj_mc + (it & 3ull), total);
      __threadfence_system();
      for (int q = 0; q < p.nranks; ++q)
        if (q != p.rank) st_release_sys(p.zflag_peer[q], it);
    }
    if (loss < p.acc && p.ctrl[1] == 0ull) p.ctrl[1] = it;
    p.ctrl[0] = it + 1ull;
  }
}

template <typename T>
static int star_iterate(const GadmmStar* p, int n_iter, cudaStream_t st) {
  const int smem = (int)sizeof(SymvSmem<T>);
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (!configured) {
    GADMM_CUDA_OK(cudaFuncSetAttribute(star_workers_kernel<T>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    GADMM_CUDA_OK(cudaFuncSetAttribute(star_server_kernel<T>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = 1;
  }
  const int nT = p->dp / kT, npairs = (nT + 1) / 2;
  const int nsp = p->nsplit > 0 ? p->nsplit : 1;
  for (int i = 0; i < n_iter; ++i) {
    star_workers_kernel<T><<<p->Nw * npairs, kSymvThreads, smem, st>>>(*p);
    star_reduce_kernel<<<nT, 256, 0, st>>>(*p);
    star_server_kernel<T><<<npairs * nsp, kSymvThreads, smem, st>>>(*p);
    star_zfinish_kernel<<<nT, 64, 0, st>>>(*p);
  }
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace gadmm

using namespace gadmm;

extern "C" int gadmm_star_iterate(const GadmmStar* p, int n_iter, void* stream) {
  GADMM_REQUIRE(p && p->dp > 0 && p->dp % kT == 0, "star: dp must be a positive multiple of 64");
  GADMM_REQUIRE(p->Nw > 0, "star: every rank must own at least one worker");
  GADMM_REQUIRE(p->nsplit >= 0 && p->nsplit <= 8 && p->zblk != nullptr,
                "star: nsplit must be in [0, 8] and zblk must be set");
  if (p->store == GADMM_STORE_F64) return star_iterate<double>(p, n_iter, (cudaStream_t)stream);
  if (p->store == GADMM_STORE_F32) return star_iterate<float>(p, n_iter, (cudaStream_t)stream);
  return set_error(-1, "star: bad store %d", p->store);
}
