// Chain GADMM / D-GADMM hot loop (K3 + K4 + K5 of SURVEY.md section 2.4) and the batched
// packed-symmetric SYMV it is built on.
//
//   gadmm_half_iter_kernel   one colour (head or tail group) of one iteration:
//       r_n = b_n - nu_n + rho*(theta_l + theta_r)      (fused into the tile producer)
//       theta_n = M_n r_n                                (streamed packed-symmetric SYMV)
//       f_n(theta_n) = 0.5 theta'(r - c rho theta) - b'theta + 0.5 y'y   (no extra matrix pass)
//       theta_n -> local replica and, over NVLink, into the replicas of the GPUs that own the
//       chain neighbours, followed by a system-scope release flag
//     (group_ADMM_closedForm.m:13-57 / :61-92; static_group_ADMM_closedForm.m:50-80)
//   gadmm_dual_objective_kernel
//       nu_n += rho(theta_n - theta_r) - rho(theta_l - theta_n)
//       all-rank objective sum through peer stores + flags, stop test, device clock, iteration++
//     (static_group_ADMM_closedForm.m:130-160; group_ADMM_closedForm.m:96-108)
//
// No NCCL and no host on this path: consumers spin on `flags` with ld.acquire.sys.
//
// Windows.  gadmm_chain_iterate() enqueues a *window* of iterations; every launch carries its
// offset inside the window and works on iteration sync[0] + offset (sync[0] is set from ctrl[0]
// by gadmm_window_begin_kernel, the window's first, fully stream-ordered launch).  That makes
// every other launch of the window -- the next iteration's head phase included -- eligible for
// programmatic dependent launch: all ordering inside a window is by data (version flags of
// theta, `nuflag` of the node duals, `sync[1]` = last iteration whose stop decision is final),
// so the head phase of iteration i+1 fills SMs while the tail phase of iteration i drains and
// while the all-rank objective exchange of iteration i is still in flight.  With the stop test
// enabled the head phase of i+1 may therefore run speculatively after the test fired at i; it
// only touches the other parity slot of theta, so the frozen state is the state of iteration i.
#include "../include/gadmm_b200.h"
#include "symv_core.cuh"

#include <stdarg.h>

namespace gadmm {

thread_local char g_last_error[512] = {0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

// ------------------------------------------------------------------------------ plain SYMV
struct PlainX {
  const double* x;
  __device__ __forceinline__ double operator()(int j) const { return x[j]; }
};

template <typename T>
__global__ void __launch_bounds__(kSymvThreads, 2)
symv_batched_kernel(const void* const* __restrict__ Sptr, const double* __restrict__ x,
                    double* __restrict__ y, int dp, double* __restrict__ scratch,
                    unsigned int* __restrict__ counters) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SymvSmem<T>& sm = *reinterpret_cast<SymvSmem<T>*>(smem_raw);
  const int nT = dp / kT, npairs = (nT + 1) / 2;
  const int k = blockIdx.x / npairs, p = blockIdx.x % npairs;
  const size_t per = (size_t)dp + colpart_elems(dp);
  double* rowpart = scratch + (size_t)k * per;
  double* colpart = rowpart + dp;
  PlainX xf{x + (size_t)k * dp};
  const bool last = symv_cta<T>(sm, reinterpret_cast<const T*>(Sptr[k]), nT, p, xf, rowpart,
                                colpart, counters + k, npairs);
  if (!last) return;
  for (int j = threadIdx.x; j < dp; j += blockDim.x)
    y[(size_t)k * dp + j] = symv_gather(rowpart, colpart, nT, j);
}

// -------------------------------------------------------------------------- chain half-iteration
struct ChainX {
  const double* b;
  const double* nu;
  const double* tl;   // neighbour replicas (may be null at the chain ends)
  const double* tr;
  double rho;
  struct Raw { double l, r, b, nu; };
  __device__ __forceinline__ Raw load(int j) const {
    Raw v;
    // a missing neighbour is decided in fin(): a select here would pull the first use of the
    // loaded value (and the wait for it) up to the load
    v.l = __ldcg((tl ? tl : nu) + j);
    v.r = __ldcg((tr ? tr : nu) + j);
    v.b = b[j];
    v.nu = __ldcg(nu + j);                       // nu is rewritten by a kernel that may overlap
    return v;
  }
  __device__ __forceinline__ double fin(const Raw& v) const {
    double t = 0.0;
    if (tl) t += v.l;
    if (tr) t += v.r;
    return fma(rho, t, v.b - v.nu);
  }
  __device__ __forceinline__ double operator()(int j) const { return fin(load(j)); }
};

// Common entry of every launch of a window: which iteration is this, and is it frozen?
//   spec = true  (head phase behind a programmatic launch): decisions through it-2 are awaited;
//                 the phase still runs when the test fired at it-1 (see the header comment)
//   spec = false: decisions through it-1 are awaited; frozen as soon as the test has fired.
// Every thread of the CTA gets the same answer (0 = nothing to do); contains a __syncthreads().
__device__ __forceinline__ unsigned long long chain_enter(const GadmmChain& c, int off, bool spec,
                                                          unsigned long long* s_it) {
  if (threadIdx.x == 0) {
    unsigned long long it = *reinterpret_cast<volatile unsigned long long*>(c.sync) +
                            (unsigned long long)off;
    bool run = it <= (unsigned long long)c.max_iter;
    if (run) {
      const unsigned long long need_dv = spec ? (it >= 2ull ? it - 2ull : 0ull) : it - 1ull;
      if (c.pdl) wait_flag_ge_gpu(c.sync + 1, need_dv);
      const unsigned long long hit = *reinterpret_cast<volatile unsigned long long*>(c.ctrl + 1);
      if (c.ctrl[2] != 0ull && hit != 0ull && hit <= need_dv) run = false;   // converged: frozen
    }
    *s_it = run ? it : 0ull;
  }
  __syncthreads();
  return *s_it;
}

// First launch of a window (ordinary stream order: everything before it has completed).
__global__ void gadmm_window_begin_kernel(const GadmmChain c) {
  const unsigned long long it = c.ctrl[0];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {
    c.sync[0] = it;
    c.sync[1] = it - 1ull;       // every earlier stop decision is final
  }
  if (i < c.Nloc) c.nuflag[i] = it - 1ull;
}

// NSTAGES = SymvCfg<T>::STAGES: two CTAs per SM (many workers per GPU).  NSTAGES doubled ("deep"):
// one CTA per SM with twice the bytes in flight, for phases with fewer active CTAs than SMs
// (one worker per GPU), where a 3-stage ring is latency-bound.
template <typename T, int NSTAGES>
__global__ void __launch_bounds__(kSymvThreads, (NSTAGES > SymvCfg<T>::STAGES) ? 1 : 2)
gadmm_half_iter_kernel(const GadmmChain c, const int color, const int off) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SymvSmem<T, NSTAGES>& sm = *reinterpret_cast<SymvSmem<T, NSTAGES>*>(smem_raw);
  __shared__ unsigned long long s_it;
  const int nT = c.dp / kT, npairs = (nT + 1) / 2, nsp = c.nsplit;
  // `order` puts workers with a neighbour on another GPU first: their theta is published early in
  // the phase and the peer's dependent CTAs (scheduled first there, too) do not sit spinning
  const int k = c.order[blockIdx.x / (npairs * nsp)], p = (blockIdx.x / nsp) % npairs,
            part = blockIdx.x % nsp;
  // let the next phase's grid start filling SMs as soon as every CTA of this one is resident
  if (c.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (c.color[k] != color) return;
  const int g = c.gid[k], l = c.left[k], r = c.right[k];
  size_t slot_in;
  {
    const unsigned long long it0 = chain_enter(c, off, c.pdl && color == 0 && off > 0, &s_it);
    if (it0 == 0ull) return;
    // heads read the tails' previous iterate, tails read this iteration's heads
    const unsigned long long need = (color == 0) ? it0 - 1 : it0;
    slot_in = (size_t)(need & 1ull) * c.Nglob;
    if (threadIdx.x == kConsumerWarps * 32) {   // producer lane 0 waits for remote neighbours
      // With programmatic dependent launch the previous phase may still be running on this GPU,
      // so local neighbours are awaited through the same version flags as remote ones.
      if (l >= 0 && (c.pdl || c.owner[l] != c.rank)) wait_flag_ge(c.flags + l, need);
      if (r >= 0 && (c.pdl || c.owner[r] != c.rank)) wait_flag_ge(c.flags + r, need);
      if (c.pdl) wait_flag_ge_gpu(c.nuflag + k, it0 - 1ull);   // own node dual of iteration it-1
    }
    __syncthreads();
  }

  ChainX xf;
  xf.b = c.b + (size_t)k * c.dp;
  xf.nu = c.nu + (size_t)k * c.dp;
  xf.tl = l >= 0 ? c.theta + (slot_in + l) * c.dp : nullptr;
  xf.tr = r >= 0 ? c.theta + (slot_in + r) * c.dp : nullptr;
  xf.rho = c.rho;

  double* rowpart = c.rowpart + (size_t)k * nsp * c.dp;
  double* colpart = c.colpart + (size_t)k * colpart_elems(c.dp);
  const bool last = symv_cta<T, ChainX, NSTAGES>(sm, reinterpret_cast<const T*>(c.Mptr[k]), nT, p,
                                                 xf, rowpart + (size_t)part * c.dp, colpart,
                                                 c.counters + k, npairs * nsp, part, nsp);
  if (!last) return;
  const unsigned long long it = *reinterpret_cast<volatile unsigned long long*>(&s_it);
  if (c.split_finish) {
    // the partial sums of worker k are complete: hand over to gadmm_finish_kernel (many CTAs)
    if (threadIdx.x == 0) {
      __threadfence();
      st_release_sys(c.wready + k, it);
    }
    return;
  }

  // ---------------------------------------------------------------- finisher (one CTA per worker)
  const size_t slot_out = (size_t)(it & 1ull) * c.Nglob;
  double* th_loc = c.theta + (slot_out + g) * c.dp;
  const int p0 = c.push0[k], p1 = c.push1[k];
  double* th_p0 = p0 >= 0 ? c.theta_peer[p0] + (slot_out + g) * c.dp : nullptr;
  double* th_p1 = p1 >= 0 ? c.theta_peer[p1] + (slot_out + g) * c.dp : nullptr;
  const double crho = c.rho * (double)c.cmul[k];
  double s1 = 0.0, s2 = 0.0;
  for (int j = threadIdx.x; j < c.dp; j += blockDim.x) {
    const double th = symv_gather(rowpart, colpart, nT, j, nsp, c.dp);
    const double rj = xf(j);
    th_loc[j] = th;
    if (th_p0) th_p0[j] = th;          // NVLink P2P store into the neighbour GPU's replica
    if (th_p1) th_p1[j] = th;
    s1 = fma(th, rj - crho * th, s1);  // theta'(G theta) with G theta = r - c rho theta
    s2 = fma(xf.b[j], th, s2);
  }
  double* red = &sm.colbuf[0][0][0];
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  if (threadIdx.x == 0) {
    c.objw[(size_t)(it & 1ull) * c.Nloc + k] = 0.5 * s1 - s2 + c.hyy[k];
    __threadfence_system();
    if (p0 >= 0) st_release_sys(c.flags_peer[p0] + g, it);
    if (p1 >= 0) st_release_sys(c.flags_peer[p1] + g, it);
    st_release_sys(c.flags + g, it);
  }
}

// ------------------------------------------------------------------ parallel finisher (large d)
// For d >= 512 the column-partial reduction is worth spreading over many CTAs: one CTA of 256
// threads per 64-entry block of theta_n.  It waits (ld.acquire) until the streaming kernel has
// published worker k's partials, adds them in a fixed order, writes theta (local replica + NVLink
// P2P stores into the neighbour GPUs), and the last block of the worker folds the per-block
// objective pieces and releases the version flags.  Launched with programmatic dependent launch
// right behind the streaming kernel, so its CTAs fill SM slots as that kernel drains.
__global__ void __launch_bounds__(256)
gadmm_finish_kernel(const GadmmChain c, const int color, const int off) {
  __shared__ double red[8];
  __shared__ double gpart[4][kT];
  __shared__ unsigned int ticket;
  __shared__ unsigned long long s_it;
  // 64 * ng threads: column tid & 63; the ng groups (tid >> 6) split the tile-rows I of the column-
  // partial sum between them.  ng = 4 for wide models (at d = 16384 a column has up to 255
  // partials: one thread adding them all is ~25 us of dependent L2 round trips on the critical
  // path of every phase; one worker per GPU: 0.521 -> 0.421 ms per iteration), ng = 1 otherwise
  // (at d = 4096 the extra threads cost more than they save: 1.314 vs 1.330 ms, same-box A/B).
  const int J = blockIdx.x, k = c.order[blockIdx.y], tid = threadIdx.x, col = tid & 63, grp = tid >> 6;
  const int ng = blockDim.x >> 6;
  if (c.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (c.color[k] != color) return;
  const unsigned long long it = chain_enter(c, off, c.pdl && color == 0 && off > 0, &s_it);
  if (it == 0ull) return;
  const int nT = c.dp / kT;
  const int g = c.gid[k], l = c.left[k], r = c.right[k];
  const unsigned long long need = (color == 0) ? it - 1 : it;
  const size_t slot_in = (size_t)(need & 1ull) * c.Nglob;
  if (tid == 0) wait_flag_ge(c.wready + k, it);
  __syncthreads();
  ChainX xf;
  xf.b = c.b + (size_t)k * c.dp;
  xf.nu = c.nu + (size_t)k * c.dp;
  xf.tl = l >= 0 ? c.theta + (slot_in + l) * c.dp : nullptr;
  xf.tr = r >= 0 ? c.theta + (slot_in + r) * c.dp : nullptr;
  xf.rho = c.rho;
  const double* rowpart = c.rowpart + (size_t)k * c.nsplit * c.dp;
  const double* colpart = c.colpart + (size_t)k * colpart_elems(c.dp);
  const size_t slot_out = (size_t)(it & 1ull) * c.Nglob;
  const int p0 = c.push0[k], p1 = c.push1[k];
  const int j = J * kT + col;
  {
    // fixed partition (I = J+1+grp, +ng, ...) and fixed combination order: deterministic
    double v = 0.0;
    if (grp == 0)
      for (int q = 0; q < c.nsplit; ++q) v += __ldcg(rowpart + (size_t)q * c.dp + j);
#pragma unroll 8
    for (int I = J + 1 + grp; I < nT; I += ng) v += __ldcg(colpart + colpart_off(I, J) + col);
    gpart[grp][col] = v;
  }
  __syncthreads();
  double s1 = 0.0, s2 = 0.0;
  if (grp == 0) {
    double th = gpart[0][col];
    for (int q = 1; q < ng; ++q) th += gpart[q][col];
    const double rj = xf(j);
    c.theta[(slot_out + g) * c.dp + j] = th;
    if (p0 >= 0) c.theta_peer[p0][(slot_out + g) * c.dp + j] = th;   // NVLink P2P stores
    if (p1 >= 0) c.theta_peer[p1][(slot_out + g) * c.dp + j] = th;
    const double crho = c.rho * (double)c.cmul[k];
    s1 = warp_sum(th * (rj - crho * th));
    s2 = warp_sum(xf.b[j] * th);
    if ((tid & 31) == 0) { red[(tid >> 5) * 2] = s1; red[(tid >> 5) * 2 + 1] = s2; }
  }
  __syncthreads();
  if (tid == 0) {
    double* sb = c.sblk + ((size_t)k * nT + J) * 2;
    sb[0] = red[0] + red[2];
    sb[1] = red[1] + red[3];
    __threadfence_system();                      // theta stores (local + peers) before the ticket
    ticket = atomicAdd(c.fin_counters + k, 1u);
  }
  __syncthreads();
  if (ticket != (unsigned)(nT - 1) || grp != 0) return;
  __threadfence();
  double a1 = 0.0, a2 = 0.0;
  for (int q = tid; q < nT; q += 64) {           // fixed partition + fixed tree: deterministic
    a1 += __ldcg(c.sblk + ((size_t)k * nT + q) * 2);
    a2 += __ldcg(c.sblk + ((size_t)k * nT + q) * 2 + 1);
  }
  a1 = warp_sum(a1);
  a2 = warp_sum(a2);
  asm volatile("bar.sync 2, 64;" ::: "memory");   // (only the 64 threads of group 0 are left)
  if ((tid & 31) == 0) { red[(tid >> 5) * 2] = a1; red[(tid >> 5) * 2 + 1] = a2; }
  asm volatile("bar.sync 2, 64;" ::: "memory");
  if (tid == 0) {
    c.fin_counters[k] = 0u;
    c.objw[(size_t)(it & 1ull) * c.Nloc + k] = 0.5 * (red[0] + red[2]) - (red[1] + red[3]) + c.hyy[k];
    __threadfence_system();
    if (p0 >= 0) st_release_sys(c.flags_peer[p0] + g, it);
    if (p1 >= 0) st_release_sys(c.flags_peer[p1] + g, it);
    st_release_sys(c.flags + g, it);
  }
}

// --------------------------------------------------------------------- dual update + objective
__global__ void __launch_bounds__(256)
gadmm_dual_objective_kernel(const GadmmChain c, const int off) {
  __shared__ double red[32];
  __shared__ unsigned int ticket;
  __shared__ unsigned long long s_it;
  if (c.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const unsigned long long it = chain_enter(c, off, false, &s_it);
  if (it == 0ull) {
    // Frozen after the stop test fired: nothing to compute, but later launches of the window
    // still wait for "the decision of iteration it-1 is final" -- keep that counter moving
    // (chain_enter has already waited for it-1, so the sequence stays ordered).
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      const unsigned long long raw = *reinterpret_cast<volatile unsigned long long*>(c.sync) +
                                     (unsigned long long)off;
      if (raw <= (unsigned long long)c.max_iter) st_release_gpu(c.sync + 1, raw);
    }
    return;
  }
  const int k = c.order[blockIdx.x];
  const int g = c.gid[k], l = c.left[k], r = c.right[k];
  if (threadIdx.x == 0) {
    if (c.pdl) wait_flag_ge(c.flags + g, it);          // own finisher (tail phase may be running)
    if (l >= 0 && (c.pdl || c.owner[l] != c.rank)) wait_flag_ge(c.flags + l, it);
    if (r >= 0 && (c.pdl || c.owner[r] != c.rank)) wait_flag_ge(c.flags + r, it);
  }
  __syncthreads();
  const size_t slot = (size_t)(it & 1ull) * c.Nglob;
  const double* th = c.theta + (slot + g) * c.dp;
  const double* tl = l >= 0 ? c.theta + (slot + l) * c.dp : nullptr;
  const double* tr = r >= 0 ? c.theta + (slot + r) * c.dp : nullptr;
  double* nu = c.nu + (size_t)k * c.dp;
  for (int j = threadIdx.x; j < c.dp; j += blockDim.x) {
    const double t = __ldcg(th + j);
    double upd = 0.0;
    if (tr) upd = upd + c.rho * (t - __ldcg(tr + j));
    if (tl) upd = upd - c.rho * (__ldcg(tl + j) - t);
    nu[j] += upd;
  }
  // ---- nu_k of iteration `it` is in place: worker k's next solve may assemble its RHS
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    st_release_gpu(c.nuflag + k, it);
    // two tickets, used alternately: the kernels of consecutive iterations may overlap
    ticket = atomicAdd(c.counters + c.Nloc + (int)(it & 1ull), 1u);
  }
  __syncthreads();
  if (ticket != (unsigned)(c.Nloc - 1)) return;
  // ---- last CTA: objective, stop test, clock
  __threadfence();
  double part = 0.0;
  const double* objw = c.objw + (size_t)(it & 1ull) * c.Nloc;
  for (int i = threadIdx.x; i < c.Nloc; i += blockDim.x) part += __ldcg(objw + i);
  part = block_sum(part, red);
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  if (lane == 0) {
    c.counters[c.Nloc + (int)(it & 1ull)] = 0u;
    // decisions are taken in iteration order even when the kernels of two iterations overlap
    if (c.pdl) wait_flag_ge_gpu(c.sync + 1, it - 1ull);
  }
  __syncwarp();
  double total = part;
  if (c.nranks > 1) {
    // all-rank sum: lane q talks to rank q (store + flag out, flag + load in), so the 2*nranks
    // NVLink round trips overlap instead of queueing behind one thread
    const int ring = (int)(it & 3ull) * c.nranks;
    double v = 0.0;
    if (lane < c.nranks) {
      *reinterpret_cast<volatile double*>(c.objpart_peer[lane] + ring + c.rank) = part;
      __threadfence_system();
      st_release_sys(c.objflag_peer[lane] + ring + c.rank, it);
      wait_flag_ge(c.objflag + ring + lane, it);
      v = *reinterpret_cast<volatile double*>(c.objpart + ring + lane);
    }
    total = 0.0;
    for (int q = 0; q < c.nranks; ++q)          // same order on every rank -> identical decision
      total += __shfl_sync(0xffffffffu, v, q);
  }
  if (lane == 0) {
    const double loss = fabs(total - c.obj0);
    c.obj_hist[it] = total;
    c.loss_hist[it] = loss;
    c.t_hist[it] = globaltimer_ns();
    if (loss < c.acc && c.ctrl[1] == 0ull) c.ctrl[1] = it;
    c.ctrl[0] = it + 1ull;
    __threadfence();
    st_release_gpu(c.sync + 1, it);             // the decision of iteration `it` is final
  }
}

template <typename T, int NSTAGES>
static int configure_half() {
  GADMM_CUDA_OK(cudaFuncSetAttribute(gadmm_half_iter_kernel<T, NSTAGES>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(SymvSmem<T, NSTAGES>)));
  return 0;
}

// `dependent`: launch with programmatic stream serialisation (the kernel may start while its
// predecessor in the stream is still running; it orders itself by flags).
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                            cudaStream_t st, bool dependent, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = dependent ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}

template <typename T, int NSTAGES>
static int launch_half_impl(const GadmmChain* c, int color, int off, cudaStream_t st) {
  const int smem = (int)sizeof(SymvSmem<T, NSTAGES>);
  const int nT = c->dp / kT, npairs = (nT + 1) / 2;
  // the head phase of a window's first iteration follows gadmm_window_begin_kernel in plain
  // stream order; every other phase is a dependent launch ordered by flags
  const bool dependent = c->pdl && !(color == 0 && off == 0);
  GADMM_CUDA_OK(launch_k(gadmm_half_iter_kernel<T, NSTAGES>, dim3(c->Nloc * npairs * c->nsplit),
                         dim3(kSymvThreads), smem, st, dependent, *c, color, off));
  return 0;
}

template <typename T>
static int launch_half(const GadmmChain* c, int color, int off, cudaStream_t st) {
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  constexpr int S1 = SymvCfg<T>::STAGES, S2 = 2 * SymvCfg<T>::STAGES;
  if (!configured) {
    if (int e = configure_half<T, S1>()) return e;
    if (int e = configure_half<T, S2>()) return e;
    configured = 1;
  }
  const int nT = c->dp / kT, npairs = (nT + 1) / 2;
  // about half of the local workers are active per phase
  const bool deep = c->nsplit == 1 && ((long long)c->Nloc * npairs + 1) / 2 <= 160;
  if (int e = deep ? launch_half_impl<T, S2>(c, color, off, st)
                   : launch_half_impl<T, S1>(c, color, off, st))
    return e;
  if (c->split_finish)
    GADMM_CUDA_OK(launch_k(gadmm_finish_kernel, dim3(nT, c->Nloc), dim3(nT >= 96 ? 256 : 64), 0, st,
                           c->pdl != 0,
                           *c, color, off));
  return 0;
}

static int chain_check(const GadmmChain* c) {
  GADMM_REQUIRE(c && c->dp > 0 && c->dp % kT == 0, "chain: dp must be a positive multiple of 64");
  GADMM_REQUIRE(c->nsplit >= 1 && c->nsplit <= 8, "chain: nsplit must be in [1, 8]");
  GADMM_REQUIRE(c->Nloc > 0, "chain: every rank must own at least one worker");
  GADMM_REQUIRE(c->nranks >= 1 && c->nranks <= 32, "chain: at most 32 ranks");
  GADMM_REQUIRE(c->sync && c->nuflag, "chain: sync / nuflag buffers missing");
  return 0;
}

static int window_begin(const GadmmChain* c, cudaStream_t st) {
  gadmm_window_begin_kernel<<<(c->Nloc + 255) / 256, 256, 0, st>>>(*c);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

static int half_iter(const GadmmChain* c, int color, int off, cudaStream_t st) {
  return c->store == GADMM_STORE_F64 ? launch_half<double>(c, color, off, st)
                                     : launch_half<float>(c, color, off, st);
}

static int dual_objective(const GadmmChain* c, int off, cudaStream_t st) {
  GADMM_CUDA_OK(launch_k(gadmm_dual_objective_kernel, dim3(c->Nloc), dim3(256), 0, st,
                         c->pdl != 0, *c, off));
  return 0;
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

const char* gadmm_last_error(void) { return g_last_error; }
int gadmm_version(void) { return 100; }

size_t gadmm_packed_elems(int dp) {
  const size_t nT = dp / kT;
  return nT * (nT + 1) / 2 * (size_t)(kT * kT);
}

size_t gadmm_symv_scratch_elems(int dp) { return (size_t)dp + colpart_elems(dp); }

int gadmm_symv_batched(const void* const* Sptr, const double* x, double* y, int N, int dp,
                       int store, double* scratch, unsigned int* counters, void* stream) {
  GADMM_REQUIRE(dp > 0 && dp % kT == 0, "symv: dp=%d must be a positive multiple of 64", dp);
  GADMM_REQUIRE(store == GADMM_STORE_F64 || store == GADMM_STORE_F32, "symv: bad store %d", store);
  if (N <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int nT = dp / kT, npairs = (nT + 1) / 2;
  if (store == GADMM_STORE_F64) {
    const int smem = (int)sizeof(SymvSmem<double>);
    GADMM_CUDA_OK(cudaFuncSetAttribute(symv_batched_kernel<double>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    symv_batched_kernel<double><<<N * npairs, kSymvThreads, smem, st>>>(Sptr, x, y, dp, scratch,
                                                                       counters);
  } else {
    const int smem = (int)sizeof(SymvSmem<float>);
    GADMM_CUDA_OK(cudaFuncSetAttribute(symv_batched_kernel<float>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    symv_batched_kernel<float><<<N * npairs, kSymvThreads, smem, st>>>(Sptr, x, y, dp, scratch,
                                                                      counters);
  }
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

/* The two single-launch entry points form a window of ONE iteration: call the head colour, the
 * tail colour and the dual kernel in that order. */
int gadmm_chain_half_iter(const GadmmChain* c, int color, void* stream) {
  if (int e = chain_check(c)) return e;
  GADMM_REQUIRE(color == 0 || color == 1, "chain: color must be 0 or 1");
  if (color == 0)
    if (int e = window_begin(c, (cudaStream_t)stream)) return e;
  return half_iter(c, color, 0, (cudaStream_t)stream);
}

int gadmm_chain_dual_objective(const GadmmChain* c, void* stream) {
  if (int e = chain_check(c)) return e;
  return dual_objective(c, 0, (cudaStream_t)stream);
}

int gadmm_chain_iterate(const GadmmChain* c, int n_iter, void* stream) {
  if (int e = chain_check(c)) return e;
  if (n_iter <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (int e = window_begin(c, st)) return e;
  for (int i = 0; i < n_iter; ++i) {
    int e;
    if ((e = half_iter(c, 0, i, st))) return e;
    if ((e = half_iter(c, 1, i, st))) return e;
    if ((e = dual_objective(c, i, st))) return e;
  }
  return 0;
}

int gadmm_chain_launches_per_iter(void) { return 3; }   /* +2 finisher launches with split_finish,
                                                             +1 per window (window begin) */

int gadmm_chain_configure(void) {
  if (int e = configure_half<double, SymvCfg<double>::STAGES>()) return e;
  if (int e = configure_half<double, 2 * SymvCfg<double>::STAGES>()) return e;
  if (int e = configure_half<float, SymvCfg<float>::STAGES>()) return e;
  return configure_half<float, 2 * SymvCfg<float>::STAGES>();
}

}  // extern "C"

// ============================================================================================
// Allocation-free batched SYMV.
//
// Who this is for: a caller that is not allowed to (or cannot) obtain device memory at the time
// it needs y = S x.  Two typical cases: (1) a plug-in living inside a host application that took
// all device memory into its own arena at start-up and forbids cudaMalloc afterwards (the call
// would fail or break the host's accounting); (2) a latency-critical control loop, where
// cudaMalloc / cudaFree are unacceptable because they may synchronise the whole device.  Such a
// caller still owns its inputs and outputs, but has nowhere to put the row/column partial sums
// and the per-matrix ticket counters that gadmm_symv_batched() expects in `scratch`/`counters`.
//
// Here that working memory is a fixed array inside the library image.  It is reserved on every
// GPU that loads the library, for the life of the process, so it is kept small:
//   GADMM_NOALLOC_MAX_MATRICES matrices per call, and
//   N * gadmm_symv_scratch_elems(dp) <= GADMM_NOALLOC_SCRATCH_DOUBLES (1 Mi doubles = 8 MiB).
// Larger requests are refused with an error message; nothing is allocated as a fallback.
//
// Sharing rules: every call on a device uses the same array.  The ticket counters are cleared at
// the start of each call (on the caller's stream); the partial sums need no clearing because
// every slot is written before it is read.  Two calls whose kernels can overlap on the GPU
// (different streams without an ordering between them) race on the array and give wrong
// results: keep all calls on one stream or order them with events.
//
// First use on a device resolves the device address of the array (cudaGetSymbolAddress) and sets
// the kernels' shared-memory attribute.  That first call does the set-up work, so make it before
// starting any stream capture on that device.
// ============================================================================================
namespace gadmm {

constexpr size_t kNoAllocMaxMatrices = GADMM_NOALLOC_MAX_MATRICES;
constexpr size_t kNoAllocCounterDoubles = kNoAllocMaxMatrices * sizeof(unsigned int) / sizeof(double);
constexpr size_t kNoAllocScratchDoubles = GADMM_NOALLOC_SCRATCH_DOUBLES;

// [ counters (kNoAllocMaxMatrices x u32) | rowpart/colpart partial sums ]
static __device__ __align__(256) double g_noalloc_arena[kNoAllocCounterDoubles + kNoAllocScratchDoubles];

constexpr int kMaxDevices = 64;
static double* g_noalloc_ptr[kMaxDevices] = {nullptr};   // device address of the arena, per device

// Once per device, with that device current.  In host code `g_noalloc_arena` is only a handle for
// the runtime -- handing it to a kernel as if it were a pointer would make the kernel touch the
// wrong memory -- so the real device address is looked up here and cached.
static int noalloc_setup(int dev, double** out) {
  if (g_noalloc_ptr[dev] == nullptr) {
    double* p = nullptr;
    GADMM_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_noalloc_arena));
    GADMM_REQUIRE(p != nullptr, "noalloc symv: cudaGetSymbolAddress returned a null address");
    GADMM_CUDA_OK(cudaFuncSetAttribute(symv_batched_kernel<double>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(SymvSmem<double>)));
    GADMM_CUDA_OK(cudaFuncSetAttribute(symv_batched_kernel<float>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(SymvSmem<float>)));
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

}  // namespace gadmm

extern "C" {

size_t gadmm_symv_noalloc_max_scratch_elems(void) { return kNoAllocScratchDoubles; }
int gadmm_symv_noalloc_max_matrices(void) { return (int)kNoAllocMaxMatrices; }

int gadmm_symv_batched_noalloc(const void* const* Sptr, const double* x, double* y, int N, int dp,
                               int store, void* stream) {
  GADMM_REQUIRE(dp > 0 && dp % kT == 0, "noalloc symv: dp=%d must be a positive multiple of 64", dp);
  GADMM_REQUIRE(store == GADMM_STORE_F64 || store == GADMM_STORE_F32,
                "noalloc symv: bad store %d", store);
  if (N <= 0) return 0;
  GADMM_REQUIRE((size_t)N <= kNoAllocMaxMatrices,
                "noalloc symv: N=%d matrices exceeds the built-in limit of %zu per call", N,
                kNoAllocMaxMatrices);
  const size_t need = (size_t)N * gadmm_symv_scratch_elems(dp);
  GADMM_REQUIRE(need <= kNoAllocScratchDoubles,
                "noalloc symv: N=%d, dp=%d needs %zu doubles of scratch but the built-in arena "
                "holds %zu; use gadmm_symv_batched with caller-provided scratch instead",
                N, dp, need, kNoAllocScratchDoubles);
  int dev = 0;
  GADMM_CUDA_OK(cudaGetDevice(&dev));
  GADMM_REQUIRE(dev >= 0 && dev < kMaxDevices, "noalloc symv: device ordinal %d out of range", dev);
  double* arena = nullptr;
  if (int e = noalloc_setup(dev, &arena)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* counters = reinterpret_cast<unsigned int*>(arena);
  double* scratch = arena + kNoAllocCounterDoubles;
  // the arena keeps whatever the previous call (or an aborted one) left: tickets must start at 0
  GADMM_CUDA_OK(cudaMemsetAsync(counters, 0, (size_t)N * sizeof(unsigned int), st));
  const int nT = dp / kT, npairs = (nT + 1) / 2;
  if (store == GADMM_STORE_F64)
    symv_batched_kernel<double><<<N * npairs, kSymvThreads, (int)sizeof(SymvSmem<double>), st>>>(
        Sptr, x, y, dp, scratch, counters);
  else
    symv_batched_kernel<float><<<N * npairs, kSymvThreads, (int)sizeof(SymvSmem<float>), st>>>(
        Sptr, x, y, dp, scratch, counters);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
