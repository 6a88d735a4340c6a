// D-GADMM re-chaining on the device (K10 of SURVEY.md section 2.4: "rechain_gather").
//
// dynamic_group_ADMM_closedForm.m:18-21 replaces the chain every tau iterations
// (`i>1 && mod(i,tau)==0`); theta and the node duals persist.  Here the chains of all coming
// windows sit in device memory and two tiny kernels, captured in the same CUDA graph as the
// iteration kernels, switch topology without the host:
//   rechain_pos_kernel     pos[path[i]] = i for the chain of the current window
//   rechain_tables_kernel  one CTA per local worker: new (left, right, colour, c_n, push targets,
//                          inverse pointer for c_n) and the fused peer gather -- theta (latest
//                          version) of neighbours that live on another GPU is pulled straight out
//                          of the owner's replica with NVLink P2P loads and its version flag is
//                          raised.  All ranks are iteration-aligned here (the per-iteration
//                          objective exchange bounds the skew to < 1 iteration), so the owners'
//                          rows are final and will not be overwritten before the next exchange.
// Both kernels are no-ops on iterations that do not re-chain.
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

__device__ __forceinline__ long long rechain_window(const GadmmRechain& p, unsigned long long it) {
  if (p.tau <= 0 || it <= 1ull || (it % (unsigned long long)p.tau) != 0ull) return -1;
  // dyn[0] = window index of paths[0], dyn[1] = number of valid rows (device-resident so that a
  // captured graph keeps working when the host uploads the next chunk of chains)
  long long kk = (long long)(it / (unsigned long long)p.tau) - (p.tau == 1 ? 1 : 0) - p.dyn[0];
  if (kk < 0) return -1;
  return kk < p.dyn[1] ? kk : p.dyn[1] - 1;
}

__global__ void rechain_pos_kernel(const GadmmRechain p) {
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  const long long kk = rechain_window(p, it);
  if (kk < 0) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < p.N) p.pos[p.paths[kk * p.N + i]] = i;
}

__global__ void __launch_bounds__(256)
rechain_tables_kernel(const GadmmRechain p) {
  const unsigned long long it = p.ctrl[0];
  if (p.ctrl[2] != 0ull && p.ctrl[1] != 0ull) return;
  const long long kk = rechain_window(p, it);
  if (kk < 0) return;
  const int k = blockIdx.x, Nl = p.Nloc;
  const int* path = p.paths + kk * p.N;
  const int g = p.tab[k];                       // row 0 of the table block: global ids
  const int ps = p.pos[g];
  const int l = ps > 0 ? path[ps - 1] : -1;
  const int r = ps < p.N - 1 ? path[ps + 1] : -1;
  if (threadIdx.x == 0) {
    const int c = (ps == 0 || ps == p.N - 1) ? 1 : 2;
    int t0 = -1, t1 = -1;
    if (l >= 0 && p.owner[l] != p.rank) t0 = p.owner[l];
    if (r >= 0 && p.owner[r] != p.rank && p.owner[r] != t0) { if (t0 < 0) t0 = p.owner[r]; else t1 = p.owner[r]; }
    p.tab[1 * Nl + k] = l;
    p.tab[2 * Nl + k] = r;
    p.tab[3 * Nl + k] = ps & 1;
    p.tab[4 * Nl + k] = c;
    p.tab[5 * Nl + k] = t0;
    p.tab[6 * Nl + k] = t1;
    p.Mptr[k] = (c == 1 ? p.base_c1 : p.base_c2) + (long long)k * p.stride_bytes;
    // Launch order of the local workers = their order along the NEW chain.  Every GPU does the
    // same, so each phase advances as one wavefront along the chain: a tail worker early in the
    // launch order depends on head workers that are early in their GPUs' launch order too.
    // (Otherwise CTAs sit spinning on flags whose producers are queued behind them elsewhere.)
    if (p.order) {
      int rank = 0;
      for (int j = 0; j < Nl; ++j) {
        const int pj = p.pos[p.tab[j]];
        rank += (pj < ps) ? 1 : 0;
      }
      p.order[rank] = k;
    }
  }
  // fused peer gather of the new remote neighbours (version it-1 lives in slot (it-1)&1)
  const unsigned long long ver = it - 1ull;
  const size_t slot = (size_t)(ver & 1ull) * p.N;
  for (int side = 0; side < 2; ++side) {
    const int nb = side == 0 ? l : r;
    if (nb < 0 || p.owner[nb] == p.rank) continue;
    const double* src = p.theta_peer[p.owner[nb]] + (slot + nb) * p.dp;
    double* dst = p.theta + (slot + nb) * p.dp;
    for (int j = threadIdx.x; j < p.dp; j += blockDim.x) dst[j] = src[j];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (l >= 0 && p.owner[l] != p.rank) atomicMax(p.flags + l, ver);
    if (r >= 0 && p.owner[r] != p.rank) atomicMax(p.flags + r, ver);
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" int gadmm_rechain(const GadmmRechain* p, void* stream) {
  GADMM_REQUIRE(p && p->N > 0 && p->Nloc > 0 && p->K > 0, "rechain: bad descriptor");
  cudaStream_t st = (cudaStream_t)stream;
  rechain_pos_kernel<<<(p->N + 255) / 256, 256, 0, st>>>(*p);
  rechain_tables_kernel<<<p->Nloc, 256, 0, st>>>(*p);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}
