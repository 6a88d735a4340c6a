// Gram / weighted Gram on the 5th-generation tensor cores (K1 of SURVEY.md section 2.4):
//     G_n += X_n' diag(w) X_n        (w = 1: X'X of group_ADMM_closedForm.m:43; w = sigma(1-sigma):
//                                     the logistic Hessian X'DX of the Newton local solve)
//
// tcgen05 has no fp64 kind, so every fp32 value is split x = hi + lo with hi, lo exactly
// representable in tf32 and the product is formed as hi*hi' + hi*lo' + lo*hi' (3 MMAs of
// kind::tf32 into one fp32 TMEM accumulator, error ~2^-21 per product).  To keep the fp32
// accumulation short, the sample dimension is cut into segments of KSEG samples; each CTA owns
// one (128x128 output tile, segment).  The segments of a tile are summed in 64-bit FIXED POINT
// (integer atomics: exact, hence independent of the order in which the CTAs arrive, so the
// result is bit-reproducible like every other kernel of the library); the scale is a power of two
// derived from s * max|x|^2 >= |G_ij|, which leaves > 40 fractional bits against the ~21 good
// bits of a split-tf32 product.  Only the lower triangle is accumulated; a last kernel converts
// to fp64, adds into G and mirrors.
//
// Data path: a prep kernel writes the (optionally sqrt(w)-scaled) shard transposed and split,
// Xt_hi / Xt_lo [N][dT][sP] fp32 with samples contiguous (K-major for both MMA operands, since
// A and B are row blocks of the same matrix).  The main kernel is warp specialised:
//   warp 0  TMA producer   cp.async.bulk.tensor.3d (box 32 x 128, SWIZZLE_128B) -> 3-stage ring
//   warp 1  MMA issuer     one elected lane: 4 k-steps x 3 split products per stage,
//                          tcgen05.commit frees the stage / signals the epilogue; owns TMEM alloc
//   warps 2-5 epilogue     tcgen05.ld 32x32b (each warp its TMEM lane quarter) -> fixed-point
//                          integer atomics on the lower triangle
// Diagonal tiles load only the A operand (B descriptor aliases it).
#include "../include/gadmm_b200.h"
#include "common.cuh"

#include <cuda.h>
#include <cstdlib>

namespace gadmm {

constexpr int GT = 128;        // output tile edge (UMMA M = N = 128)
constexpr int GBK = 32;        // fp32 elements per k-chunk = one 128-byte swizzle row
constexpr int GSTAGES = 3;
constexpr int GTILE_BYTES = GT * GBK * 4;        // 16 KiB per operand tile
constexpr int GSTAGE_BYTES = 4 * GTILE_BYTES;    // A_hi, A_lo, B_hi, B_lo
constexpr int GTHREADS = 192;
constexpr int GSMEM = GSTAGES * GSTAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;

__device__ __forceinline__ float tf32_round(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}

// X: [N][s][d] (fp64 or fp32) -> Xt_hi, Xt_lo: [N][dT][sP] fp32, zero padded, optionally scaled by
// sqrt(w[n][r]).  32x32 smem transpose: coalesced reads along d, coalesced writes along s.
template <typename TIn>
__global__ void __launch_bounds__(256)
split_transpose_kernel(const TIn* __restrict__ X, const double* __restrict__ w, int s, int d, int sP,
                       int dT, float* __restrict__ hi, float* __restrict__ lo,
                       unsigned int* __restrict__ xmax) {
  __shared__ float tile[32][33];
  float vmax = 0.f;
  const int n = blockIdx.z, r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const TIn* Xn = X + (size_t)n * s * d;
  for (int i = ty; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + tx;
    float v = 0.f;
    if (r < s && c < d) {
      double x = (double)Xn[(size_t)r * d + c];
      if (w) x *= sqrt(w[(size_t)n * s + r]);
      v = (float)x;
    }
    tile[i][tx] = v;
    vmax = fmaxf(vmax, fabsf(v));
  }
  // max |x| of matrix n (non-negative floats order like their bit patterns: exact, order-free)
  for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
  __shared__ float wmax[8];
  if (tx == 0) wmax[ty] = vmax;
  __syncthreads();
  if (threadIdx.x == 0) {                  // one atomic per block, and only if it can raise the max
    for (int q = 1; q < 8; ++q) vmax = fmaxf(vmax, wmax[q]);
    const unsigned int bits = __float_as_uint(vmax);
    if (bits > *reinterpret_cast<volatile unsigned int*>(xmax + n)) atomicMax(xmax + n, bits);
  }
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + tx;      // output row = feature c, column = sample r
    if (c < dT && r < sP) {
      const float v = tile[tx][i];
      const float h = tf32_round(v);
      const size_t o = ((size_t)n * dT + c) * sP + r;
      hi[o] = h;
      lo[o] = tf32_round(v - h);
    }
  }
}

// scale[n] = 2^e with s * xmax^2 * 2^e < 2^61: every partial sum of |G_ij| <= s xmax^2 fits.
__global__ void gram_scale_kernel(const unsigned int* __restrict__ xmax, double* __restrict__ scale,
                                  int N, int sP) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const double xm = (double)__uint_as_float(xmax[n]);
  const double bound = fmax((double)sP * xm * xm, 1e-300);
  int e;
  frexp(bound, &e);                       // bound < 2^e
  scale[n] = ldexp(1.0, 60 - e);
}

__device__ __forceinline__ void fixed_add(long long* acc, double val, double scale) {
  atomicAdd(reinterpret_cast<unsigned long long*>(acc),
            static_cast<unsigned long long>(__double2ll_rn(val * scale)));
}

// G (+)= acc / scale: 32x32 tiles of the lower triangle; the mirror image goes through a shared
// memory transpose so that both the lower and the upper tile are updated with coalesced rows.
__global__ void __launch_bounds__(256)
gram_fixed_finish_kernel(const long long* __restrict__ acc, const double* __restrict__ scale,
                         double* __restrict__ G, int dp) {
  __shared__ double tile[32][33];
  const int n = blockIdx.z, ti = blockIdx.y, tj = blockIdx.x;
  if (tj > ti) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const double inv = 1.0 / scale[n];                        // power of two: exact
  const long long* An = acc + (size_t)n * dp * dp;
  double* Gn = G + (size_t)n * dp * dp;
  for (int r = ty; r < 32; r += 8) {
    const int i = ti * 32 + r, j = tj * 32 + tx;
    double g = 0.0;
    if (i < dp && j < dp && j <= i) {
      g = (double)An[(size_t)i * dp + j] * inv;
      Gn[(size_t)i * dp + j] += g;
    }
    tile[r][tx] = g;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {                        // upper image: element (j, i) = (i, j)
    const int j = tj * 32 + r, i = ti * 32 + tx;
    if (i < dp && j < dp && j < i) Gn[(size_t)j * dp + i] += tile[tx][r];
  }
}

// ------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0,
                                            int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, kind::tf32, fp32 accumulate
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, SWIZZLE_128B operand tile: rows of 128 B, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(const void* smem) {
  const uint32_t a = smem_u32(smem);
  uint64_t d = 0;
  d |= (uint64_t)((a >> 4) & 0x3FFF);              // start address  [0,14)
  d |= (uint64_t)1 << 16;                          // leading byte offset (unused for SW128 K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                // stride byte offset [32,46)
  d |= (uint64_t)1 << 46;                          // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                          // SWIZZLE_128B
  return d;
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// idesc: c=F32 (1<<4), a=b=TF32 (2<<7, 2<<10), K-major both, N>>3 at [17,23), M>>4 at [24,29)
constexpr uint32_t kIdescTf32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(GT >> 3) << 17) |
                                ((uint32_t)(GT >> 4) << 24);

// grid: (lower-triangular 128-tiles, k-segments, N)
__global__ void __launch_bounds__(GTHREADS, 1)
gram_tcgen05_kernel(const __grid_constant__ CUtensorMap map_hi,
                    const __grid_constant__ CUtensorMap map_lo, int sP, int kseg, int dp,
                    long long* __restrict__ acc, const double* __restrict__ scale,
                    double* __restrict__ G) {
  extern __shared__ unsigned char gsm_raw[];
  unsigned char* base =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(gsm_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(base + GSTAGES * GSTAGE_BYTES);
  uint64_t* empty = full + GSTAGES;
  uint64_t* acc_full = empty + GSTAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int I = 0, rem = blockIdx.x;
  while (rem > I) { rem -= I + 1; ++I; }
  const int J = rem, n = blockIdx.z;
  const bool diag = (I == J);
  const int k_begin = blockIdx.y * kseg;
  const int k_end = min(sP, k_begin + kseg);
  const int nchunks = (k_end - k_begin) / GBK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < GSTAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) {   // TMEM: 128 fp32 accumulator columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"(128)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int c = 0; c < nchunks; ++c) {
        mbar_wait(&empty[stage], phase ^ 1);
        unsigned char* st = base + stage * GSTAGE_BYTES;
        const int k0 = k_begin + c * GBK;
        mbar_arrive_expect_tx(&full[stage], diag ? 2 * GTILE_BYTES : 4 * GTILE_BYTES);
        tma_load_3d(st, &map_hi, &full[stage], k0, I * GT, n);
        tma_load_3d(st + GTILE_BYTES, &map_lo, &full[stage], k0, I * GT, n);
        if (!diag) {
          tma_load_3d(st + 2 * GTILE_BYTES, &map_hi, &full[stage], k0, J * GT, n);
          tma_load_3d(st + 3 * GTILE_BYTES, &map_lo, &full[stage], k0, J * GT, n);
        }
        if (++stage == GSTAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int c = 0; c < nchunks; ++c) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        unsigned char* st = base + stage * GSTAGE_BYTES;
        const uint64_t a_hi = umma_desc_sw128(st), a_lo = umma_desc_sw128(st + GTILE_BYTES);
        const uint64_t b_hi = diag ? a_hi : umma_desc_sw128(st + 2 * GTILE_BYTES);
        const uint64_t b_lo = diag ? a_lo : umma_desc_sw128(st + 3 * GTILE_BYTES);
#pragma unroll
        for (int k = 0; k < GBK / 8; ++k) {                 // UMMA_K = 8 tf32 = 32 bytes
          const uint64_t adv = (uint64_t)((k * 32) >> 4);   // start-address field advance
          tc_mma_tf32(tmem, a_hi + adv, b_hi + adv, kIdescTf32, (c | k) != 0);
          tc_mma_tf32(tmem, a_hi + adv, b_lo + adv, kIdescTf32, 1u);
          tc_mma_tf32(tmem, a_lo + adv, b_hi + adv, kIdescTf32, 1u);
        }
        tc_commit(&empty[stage]);                            // frees the smem stage when MMAs retire
        if (++stage == GSTAGES) { stage = 0; phase ^= 1; }
      }
      tc_commit(acc_full);                                   // accumulator complete
    }
  } else {
    // ---- epilogue: warps 2..5 own TMEM lane quarters (warp id % 4)
    const int q = warp & 3;
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int m = q * 32 + lane;                             // row inside the tile
    // One K segment: this CTA is the only writer of its tile, so fp64 reductions straight into G
    // (and its mirror image) are already order-free.  Several segments: fixed point.
    const bool direct = gridDim.y == 1;
    long long* An = acc + (size_t)n * dp * dp;
    double* Gn = G + (size_t)n * dp * dp;
    const double sc = direct ? 1.0 : scale[n];
    const size_t grow = (size_t)I * GT + m;
    if (nchunks > 0) {
      for (int cb = 0; cb < GT; cb += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem + ((uint32_t)(q * 32) << 16) + cb, v);
        if (grow < (size_t)dp) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const size_t gcol = (size_t)J * GT + cb + j;
            if (gcol >= (size_t)dp) continue;
            if (diag && gcol > grow) continue;               // lower triangle of diagonal tiles
            const double val = (double)__uint_as_float(v[j]);
            if (direct) {            // single writer per element: the reduction is a plain add
              atomicAdd(Gn + grow * dp + gcol, val);        // (RED: fire-and-forget, no read stall)
              if (gcol != grow) atomicAdd(Gn + gcol * dp + grow, val);
            } else {
              fixed_add(An + grow * dp + gcol, val, sc);
            }
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128)
                 : "memory");
  }
}

// ------------------------------------------------------------------------------ CTA-pair variant
// The single-CTA kernel above moves 64 KiB of operands per 12 MMAs (768 tensor cycles): 85 B/clk
// per SM, twice what L2 delivers to all 148 SMs at once, and measures at ~51 % tensor activity.
// Here two CTAs of a cluster (one TPC) share one 256x256 output tile with cta_group::2 MMAs
// (M = 256: 128 accumulator rows per CTA, N = 256): each CTA loads only its 128 rows of A and its
// 128-row half of B, so the same 64 KiB per stage now feed 12 MMAs of twice the size.
//   both CTAs   warp 0 lane 0: TMA (cta_group::2 form, completes on the LEADER's full barrier)
//   leader      warp 1 lane 0: MMAs for the pair; commits multicast to both CTAs' empty barriers
//   both CTAs   warps 2-5: epilogue of their own 128 x 256 accumulator half
// On diagonal tiles a CTA's B half is its own A tile (descriptor aliases it, 2 loads per stage).
constexpr int GT2 = 256;
constexpr uint32_t kIdescTf32Pair = (1u << 4) | (2u << 7) | (2u << 10) |
                                    ((uint32_t)(GT2 >> 3) << 17) | ((uint32_t)(GT2 >> 4) << 24);

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                   : "memory");
}
__device__ __forceinline__ uint32_t mapa_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// TMA tile load of the executing CTA's smem, byte count reported to `bar_cluster_addr` (an
// mbarrier anywhere in the CTA pair)
__device__ __forceinline__ void tma_load_3d_pair(void* dst, const CUtensorMap* map,
                                                 uint32_t bar_cluster_addr, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {   // arrives in both CTAs
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                                 uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// grid: (2 * lower-triangular 256-tiles, k-segments, N), clusters of 2 along x
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GTHREADS, 1)
gram_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap map_hi,
                         const __grid_constant__ CUtensorMap map_lo, int sP, int kseg, int dp,
                         long long* __restrict__ acc, const double* __restrict__ scale,
                         double* __restrict__ G) {
  extern __shared__ unsigned char gsm_raw[];
  unsigned char* base =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(gsm_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(base + GSTAGES * GSTAGE_BYTES);
  uint64_t* empty = full + GSTAGES;
  uint64_t* acc_full = empty + GSTAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  int I = 0, rem = blockIdx.x >> 1;
  while (rem > I) { rem -= I + 1; ++I; }
  const int J = rem, n = blockIdx.z;
  const bool diag = (I == J);
  const int k_begin = blockIdx.y * kseg;
  const int k_end = min(sP, k_begin + kseg);
  const int nchunks = (k_end - k_begin) / GBK;
  // The tensor core's fp32 accumulation error grows with the chain length, the epilogue cost
  // shrinks with it: the first half of the segment accumulates into TMEM columns [0,256), the
  // second half into [256,512), and the epilogue adds the two in fp64.
  const int nhalf = (nchunks + 1) / 2;
  const int arow = I * GT2 + (int)rank * GT;      // this CTA's 128 rows of A (= its accumulator rows)
  const int brow = J * GT2 + (int)rank * GT;      // this CTA's half of the 256 B rows (tile columns)

  if (threadIdx.x == 0) {
    for (int s = 0; s < GSTAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) {   // two 256-column fp32 accumulators in each CTA of the pair (all of TMEM)
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_slot)),
                 "r"(2 * GT2)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers are initialised before anything targets them
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t per_cta = diag ? 2 * GTILE_BYTES : 4 * GTILE_BYTES;
      for (int c = 0; c < nchunks; ++c) {
        mbar_wait(&empty[stage], phase ^ 1);
        unsigned char* st = base + stage * GSTAGE_BYTES;
        const int k0 = k_begin + c * GBK;
        if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * per_cta);     // both CTAs' bytes
        const uint32_t lbar = mapa_cta(smem_u32(&full[stage]), 0);
        tma_load_3d_pair(st, &map_hi, lbar, k0, arow, n);
        tma_load_3d_pair(st + GTILE_BYTES, &map_lo, lbar, k0, arow, n);
        if (!diag) {
          tma_load_3d_pair(st + 2 * GTILE_BYTES, &map_hi, lbar, k0, brow, n);
          tma_load_3d_pair(st + 3 * GTILE_BYTES, &map_lo, lbar, k0, brow, n);
        }
        if (++stage == GSTAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int c = 0; c < nchunks; ++c) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        unsigned char* st = base + stage * GSTAGE_BYTES;
        const uint64_t a_hi = umma_desc_sw128(st), a_lo = umma_desc_sw128(st + GTILE_BYTES);
        const uint64_t b_hi = diag ? a_hi : umma_desc_sw128(st + 2 * GTILE_BYTES);
        const uint64_t b_lo = diag ? a_lo : umma_desc_sw128(st + 3 * GTILE_BYTES);
#pragma unroll
        for (int k = 0; k < GBK / 8; ++k) {
          const uint64_t adv = (uint64_t)((k * 32) >> 4);
          const uint32_t acc = tmem + (c >= nhalf ? GT2 : 0);
          tc_mma_tf32_pair(acc, a_hi + adv, b_hi + adv, kIdescTf32Pair,
                           (k != 0 || (c != 0 && c != nhalf)) ? 1u : 0u);
          tc_mma_tf32_pair(acc, a_hi + adv, b_lo + adv, kIdescTf32Pair, 1u);
          tc_mma_tf32_pair(acc, a_lo + adv, b_hi + adv, kIdescTf32Pair, 1u);
        }
        tc_commit_pair(&empty[stage]);
        if (++stage == GSTAGES) { stage = 0; phase ^= 1; }
      }
      tc_commit_pair(acc_full);
    }
  } else {
    const int q = warp & 3;
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int m = q * 32 + lane;
    const bool direct = gridDim.y == 1;
    long long* An = acc + (size_t)n * dp * dp;
    double* Gn = G + (size_t)n * dp * dp;
    const double sc = direct ? 1.0 : scale[n];
    const size_t grow = (size_t)arow + m;
    if (nchunks > 0) {
      for (int cb = 0; cb < GT2; cb += 32) {
        if ((size_t)J * GT2 + cb >= (size_t)dp) break;                     // warp-uniform
        if (diag && (size_t)J * GT2 + cb > (size_t)arow + GT - 1) break;   // above the diagonal
        uint32_t v[32], v2[32];
        tmem_ld_32x32b_x32(tmem + ((uint32_t)(q * 32) << 16) + cb, v);
        if (nchunks > nhalf) {
          tmem_ld_32x32b_x32(tmem + ((uint32_t)(q * 32) << 16) + GT2 + cb, v2);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v2[j] = 0u;
        }
        if (grow < (size_t)dp) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const size_t gcol = (size_t)J * GT2 + cb + j;
            if (gcol >= (size_t)dp) continue;
            if (diag && gcol > grow) continue;
            const double val = (double)__uint_as_float(v[j]) + (double)__uint_as_float(v2[j]);
            if (direct) {            // single writer per element: the reduction is a plain add
              atomicAdd(Gn + grow * dp + gcol, val);        // (RED: fire-and-forget, no read stall)
              if (gcol != grow) atomicAdd(Gn + gcol * dp + grow, val);
            } else {
              fixed_add(An + grow * dp + gcol, val, sc);
            }
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  cluster_sync_all();          // nobody leaves (or frees TMEM) while the pair's MMAs may still read
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * GT2)
                 : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int make_map(CUtensorMap* map, const float* ptr, int N, int dT, int sP) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    cudaDriverEntryPointQueryResult qres;
    void* p = nullptr;
    GADMM_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    GADMM_REQUIRE(p && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  const cuuint64_t dims[3] = {(cuuint64_t)sP, (cuuint64_t)dT, (cuuint64_t)N};
  const cuuint64_t strides[2] = {(cuuint64_t)sP * 4, (cuuint64_t)sP * dT * 4};
  const cuuint32_t box[3] = {GBK, GT, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr), dims,
                        strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  GADMM_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

/* split operands (2 x N x dT x sP floats) + fixed-point accumulators (N x dp x dp int64, dp <= dT)
 * + per-matrix max|x| and scale */
size_t gadmm_gram_tc_workspace_floats(int N, int s, int d) {
  const size_t dT = (size_t)(d + GT - 1) / GT * GT, sP = (size_t)(s + GBK - 1) / GBK * GBK;
  return 2 * (size_t)N * dT * sP + 2 * (size_t)N * dT * dT + 4 * (size_t)N + 64;
}

int gadmm_gram_tcgen05(const void* X, int x_is_f32, const double* w, int N, int s, int d, int dp,
                       float* work, double* G, int accumulate, int kseg, void* stream) {
  GADMM_REQUIRE(N > 0 && s > 0 && d > 0 && dp >= d && dp % 64 == 0, "gram_tcgen05: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const int dT = (d + GT - 1) / GT * GT, sP = (s + GBK - 1) / GBK * GBK;
  // (CTA pairs on 256x256 tiles once there are enough of them; GADMM_GRAM_PAIR=0/1 forces a path)
  GADMM_REQUIRE(dp <= dT, "gram_tcgen05: dp=%d exceeds the padded width %d", dp, dT);
  float* hi = work;
  float* lo = work + (size_t)N * dT * sP;
  // (offsets below keep 16-byte alignment: every term is a multiple of 4 floats)
  long long* acc = reinterpret_cast<long long*>(lo + (size_t)N * dT * sP);
  double* scale = reinterpret_cast<double*>(acc + (size_t)N * dT * dT);
  unsigned int* xmax = reinterpret_cast<unsigned int*>(scale + N);
  bool pair = dT >= 512;
  if (const char* e = getenv("GADMM_GRAM_PAIR")) pair = (e[0] == '1');
  if (kseg <= 0) kseg = pair ? 4096 : 2048;      // the pair's epilogue is twice as long: amortise it
  kseg = (kseg + GBK - 1) / GBK * GBK;
  const int nseg = (sP + kseg - 1) / kseg;
  if (nseg > 1) GADMM_CUDA_OK(cudaMemsetAsync(acc, 0, (size_t)N * dp * dp * sizeof(long long), st));
  GADMM_CUDA_OK(cudaMemsetAsync(xmax, 0, (size_t)N * sizeof(unsigned int), st));
  dim3 pg((sP + 31) / 32, (dT + 31) / 32, N);
  if (x_is_f32)
    split_transpose_kernel<float><<<pg, 256, 0, st>>>((const float*)X, w, s, d, sP, dT, hi, lo, xmax);
  else
    split_transpose_kernel<double><<<pg, 256, 0, st>>>((const double*)X, w, s, d, sP, dT, hi, lo, xmax);
  gram_scale_kernel<<<(N + 127) / 128, 128, 0, st>>>(xmax, scale, N, sP);
  GADMM_CUDA_OK(cudaGetLastError());
  if (!accumulate) GADMM_CUDA_OK(cudaMemsetAsync(G, 0, (size_t)N * dp * dp * sizeof(double), st));
  CUtensorMap mh, ml;
  if (int e = make_map(&mh, hi, N, dT, sP)) return e;
  if (int e = make_map(&ml, lo, N, dT, sP)) return e;
  static PerDeviceState cfg_state;
  int& configured = cfg_state.here();
  if (!configured) {
    GADMM_CUDA_OK(cudaFuncSetAttribute(gram_tcgen05_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GSMEM));
    GADMM_CUDA_OK(cudaFuncSetAttribute(gram_tcgen05_pair_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GSMEM));
    configured = 1;
  }
  if (pair) {
    const int nB2 = (dT + GT2 - 1) / GT2;
    dim3 grid(2 * (nB2 * (nB2 + 1) / 2), nseg, N);
    gram_tcgen05_pair_kernel<<<grid, GTHREADS, GSMEM, st>>>(mh, ml, sP, kseg, dp, acc, scale, G);
  } else {
    const int nB = dT / GT;
    dim3 grid(nB * (nB + 1) / 2, nseg, N);
    gram_tcgen05_kernel<<<grid, GTHREADS, GSMEM, st>>>(mh, ml, sP, kseg, dp, acc, scale, G);
  }
  if (nseg > 1)
    gram_fixed_finish_kernel<<<dim3((dp + 31) / 32, (dp + 31) / 32, N), 256, 0, st>>>(acc, scale, G, dp);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
