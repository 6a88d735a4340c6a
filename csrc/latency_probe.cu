// NVLink flag-latency probes (profiles/flag_latency.md): how long the primitives of the halo
// protocol take on this box.  No counterpart in the reference (its "messages" are array reads in
// one MATLAB process); these kernels measure the cost of the messages the product really sends.
//
//   pingpong: rank A stores a payload word + st.release.sys flag into rank B's memory, B spins
//             with ld.acquire.sys and answers the same way -> round-trip time per hop pair.
//   exchange: the all-rank scalar exchange of the dual/objective kernels -- lane q of one warp
//             stores (value, flag) into rank q and waits for rank q's flag -> time per exchange.
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

__global__ void pingpong_kernel(unsigned long long* my_flag, double* my_word,
                                unsigned long long* peer_flag, double* peer_word, int iters, int role,
                                unsigned long long* out_ns) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const unsigned long long t0 = globaltimer_ns();
  for (int i = 1; i <= iters; ++i) {
    if (role == 0) {
      *reinterpret_cast<volatile double*>(peer_word) = (double)i;
      __threadfence_system();
      st_release_sys(peer_flag, (unsigned long long)i);
      wait_flag_ge(my_flag, (unsigned long long)i);
    } else {
      wait_flag_ge(my_flag, (unsigned long long)i);
      *reinterpret_cast<volatile double*>(peer_word) = *reinterpret_cast<volatile double*>(my_word);
      __threadfence_system();
      st_release_sys(peer_flag, (unsigned long long)i);
    }
  }
  out_ns[0] = globaltimer_ns() - t0;
}

__global__ void exchange_kernel(double* part, unsigned long long* flag, double* const* part_peer,
                                unsigned long long* const* flag_peer, int rank, int nranks, int iters,
                                unsigned long long base, unsigned long long* out_ns, double* out_sum) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  const unsigned long long t0 = globaltimer_ns();
  double total = 0.0;
  for (int i = 1; i <= iters; ++i) {
    const unsigned long long it = base + (unsigned long long)i;
    const int ring = (int)(it & 3ull) * nranks;
    double v = 0.0;
    if (lane < nranks) {
      *reinterpret_cast<volatile double*>(part_peer[lane] + ring + rank) = (double)(rank + 1);
      __threadfence_system();
      st_release_sys(flag_peer[lane] + ring + rank, it);
      wait_flag_ge(flag + ring + lane, it);
      v = *reinterpret_cast<volatile double*>(part + ring + lane);
    }
    total = 0.0;
    for (int q = 0; q < nranks; ++q) total += __shfl_sync(0xffffffffu, v, q);
  }
  if (lane == 0) {
    out_ns[0] = globaltimer_ns() - t0;
    out_sum[0] = total;
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

int gadmm_probe_pingpong(unsigned long long* my_flag, double* my_word, unsigned long long* peer_flag,
                         double* peer_word, int iters, int role, unsigned long long* out_ns,
                         void* stream) {
  GADMM_REQUIRE(iters > 0 && (role == 0 || role == 1), "pingpong: bad arguments");
  pingpong_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(my_flag, my_word, peer_flag, peer_word, iters,
                                                      role, out_ns);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_probe_exchange(double* part, unsigned long long* flag, double* const* part_peer,
                         unsigned long long* const* flag_peer, int rank, int nranks, int iters,
                         unsigned long long base, unsigned long long* out_ns, double* out_sum,
                         void* stream) {
  GADMM_REQUIRE(iters > 0 && nranks >= 1 && nranks <= 32, "exchange probe: bad arguments");
  exchange_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(part, flag, part_peer, flag_peer, rank, nranks,
                                                      iters, base, out_ns, out_sum);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
