// Shared device helpers for the gadmm_b200 native library (sm_100a only).
//
// PTX wrappers for the Blackwell/Hopper async machinery used by the hot kernels:
//   * mbarrier (init / arrive / expect_tx / try_wait.parity)
//   * cp.async.bulk (1-D bulk copy global -> shared, completes on an mbarrier; SASS: UBLKCP)
//   * system-scope release/acquire flag accesses for the NVLink P2P halo protocol
//   * %globaltimer for device-side per-iteration timestamps
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define GADMM_TILE 64  // edge of one packed-symmetric tile (elements)

namespace gadmm {

// cudaFuncSetAttribute applies to the CURRENT device only: launch helpers remember per device what
// they have configured (a process that drives several GPUs configures each one on first use).
struct PerDeviceState {
  static constexpr int kMax = 64;
  int v[kMax] = {0};
  int& here() {
    int dev = 0;
    cudaGetDevice(&dev);
    return v[(dev >= 0 && dev < kMax) ? dev : kMax - 1];
  }
};


extern thread_local char g_last_error[512];
int set_error(int code, const char* fmt, ...);

#define GADMM_CUDA_OK(expr)                                                              \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess)                                                               \
      return ::gadmm::set_error((int)_e, "%s failed: %s (%s:%d)", #expr,                \
                                cudaGetErrorString(_e), __FILE__, __LINE__);             \
  } while (0)

#define GADMM_REQUIRE(cond, ...)                                                         \
  do {                                                                                   \
    if (!(cond)) return ::gadmm::set_error(-1, __VA_ARGS__);                             \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)
               : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ------------------------------------------------------------------------------ bulk async copy
// 1-D bulk copy global -> shared::cta; completion is signalled as complete_tx on `bar`.
// Both addresses 16-byte aligned, `bytes` a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// Same with an L2 eviction-priority hint (streamed-once operands: evict_first).
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                              uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// ------------------------------------------------------------------- cross-GPU flag protocol
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Same-GPU producer (another kernel or CTA of this process): device scope is enough.
__device__ __forceinline__ void wait_flag_ge_gpu(const unsigned long long* p, unsigned long long want) {
  while (ld_acquire_gpu(p) < want) {
    __nanosleep(20);
  }
}
// Spin until *p >= want.  `abort_flag` (may be null) lets the host break a wedged wait.
__device__ __forceinline__ void wait_flag_ge(const unsigned long long* p, unsigned long long want) {
  while (ld_acquire_sys(p) < want) {
    __nanosleep(32);
  }
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ------------------------------------------------------------------------------- reductions
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Deterministic block sum (fixed tree).  `red` holds >= blockDim.x/32 doubles.  All threads call.
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  double t = 0.0;
  if (w == 0) {
    for (int i = lane; i < nw; i += 32) t += red[i];
    t = warp_sum(t);
  }
  __syncthreads();
  if (threadIdx.x == 0) red[0] = t;
  __syncthreads();
  t = red[0];
  __syncthreads();
  return t;
}

}  // namespace gadmm
