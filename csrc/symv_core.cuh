// Packed-symmetric streaming SYMV core (one CTA = two tile-rows of one matrix).
//
// Layout: lower block-triangle of 64x64 row-major tiles, tile (I,J) at index I(I+1)/2+J, diagonal
// tiles stored full.  Reading only the lower triangle halves the HBM bytes of y = S x, which is
// the whole cost of a GADMM half-iteration (d^2 bytes per worker, memory bound).
//
// CTA structure (288 threads): warp 8 is the producer -- it streams tiles global -> shared with
// cp.async.bulk (UBLKCP) through a STAGES-deep mbarrier ring and assembles the matching 64-entry
// slice of x (the fused RHS) beside every tile, a group of tiles at a time with the bulk copies
// running one group ahead (see the producer loop); warps 0..7 are consumers -- warp w owns rows
// 8w..8w+7 of every tile, lane l owns columns 2l,2l+1.  Per element: one FMA into the row sum
// (y_I += T x_J) and one into the column sum (y_J += T' x_I), all in fp64, no shuffles in the
// inner loop.  Row sums are finished inside the CTA; column sums go to `colpart` and are added
// by the finisher (the last CTA of the matrix, found with an atomic ticket) in a fixed order, so
// results are bit-reproducible run to run.
#pragma once
#include <type_traits>
#include "common.cuh"

namespace gadmm {

constexpr int kT = GADMM_TILE;           // 64
constexpr int kConsumerWarps = 8;
constexpr int kSymvThreads = (kConsumerWarps + 1) * 32;  // 288
constexpr int kRowsPerWarp = kT / kConsumerWarps;        // 8
// (One bulk copy per tile: splitting a tile over several issuing lanes made no difference.)

template <typename T> struct SymvCfg;
template <> struct SymvCfg<double> { static constexpr int STAGES = 3; };
template <> struct SymvCfg<float>  { static constexpr int STAGES = 6; };

// RHS functors may split into raw loads and the arithmetic on them (`Raw`, `load`, `fin`), so that
// the producer can keep the loads of several tiles in flight; functors without the split are
// used as they are.
template <typename XF, typename = void> struct XSplit {
  using Raw = double;
  static __device__ __forceinline__ Raw load(const XF& xf, int j) { return xf(j); }
  static __device__ __forceinline__ double fin(const XF&, const Raw& r) { return r; }
};
template <typename XF> struct XSplit<XF, std::void_t<typename XF::Raw>> {
  using Raw = typename XF::Raw;
  static __device__ __forceinline__ Raw load(const XF& xf, int j) { return xf.load(j); }
  static __device__ __forceinline__ double fin(const XF& xf, const Raw& r) { return xf.fin(r); }
};

// Tiles per producer group (see the producer loop).
#ifndef GADMM_XGROUP_F32
#define GADMM_XGROUP_F32 3
#endif
#ifndef GADMM_XGROUP_F64
#define GADMM_XGROUP_F64 2
#endif
template <typename T> struct XGroup;
template <> struct XGroup<double> { static constexpr int G = GADMM_XGROUP_F64; };
template <> struct XGroup<float>  { static constexpr int G = GADMM_XGROUP_F32; };

template <typename T, int NSTAGES = SymvCfg<T>::STAGES>
struct SymvSmem {
  static constexpr int STAGES = NSTAGES;
  alignas(128) T tile[STAGES][kT * kT];
  alignas(16) double xs[STAGES][kT];          // x_J slice travelling with each tile
  alignas(16) double xrow[2][kT];             // x_I slices of the CTA's two tile-rows
  alignas(16) double colbuf[2][kConsumerWarps][kT];
  alignas(8) uint64_t full[STAGES];
  alignas(8) uint64_t empty[STAGES];
  unsigned int ticket;
};

__host__ __device__ inline size_t colpart_elems(int dp) {
  const size_t nT = dp / kT;
  return (size_t)kT * (nT * (nT - 1) / 2);   // sum over tile-rows I of I*64 column partials
}
// colpart layout: for tile-row I, J<I: 64 doubles at  (I*(I-1)/2 + J) * 64
__host__ __device__ inline size_t colpart_off(int I, int J) {
  return ((size_t)I * (I - 1) / 2 + J) * kT;
}

__device__ __forceinline__ double2 ld_pair(const double* p) {
  return *reinterpret_cast<const double2*>(p);
}
// fp32 -> fp64 widening.  F2F.F64.F32 runs on the quarter-rate XU pipe, the busiest pipe of the
// fp32-tile kernel (34 %, ncu: profiles/half_iter_f32_ncu_raw_r2.csv).  GADMM_F2D_INT=1 widens with
// integer ops on the ALU pipe instead (exact for normal numbers and zero, fp32 denormals flush to
// zero); measured 4.4 % SLOWER in a same-box A/B, so it is off: the kernel was starved by its
// producer, not bound by a pipe (profiles/half_iter_fp64.md).
#ifndef GADMM_F2D_INT
#define GADMM_F2D_INT 0
#endif
__device__ __forceinline__ double widen_f32(float f) {
#if GADMM_F2D_INT
  const unsigned int u = __float_as_uint(f);
  const unsigned int mag = u & 0x7fffffffu;
  unsigned int hi = (mag >> 3) + 0x38000000u;            // exponent re-biased by 1023 - 127 = 896
  unsigned int lo = u << 29;
  if (mag < 0x00800000u) { hi = 0u; lo = 0u; }           // zero / denormal
  return __hiloint2double((int)(hi | (u & 0x80000000u)), (int)lo);
#else
  return (double)f;
#endif
}
__device__ __forceinline__ double2 ld_pair(const float* p) {
  const float2 f = *reinterpret_cast<const float2*>(p);
  return make_double2(widen_f32(f.x), widen_f32(f.y));
}

// Streams the CTA's tiles and writes rowpart / colpart.  Returns true (to every thread) in the
// CTA that completed matrix `k` last.  `xf(j)` yields element j of the input vector.
//   S        packed matrix of this worker
//   pairIdx  which pair of tile-rows (pairIdx, nT-1-pairIdx) this CTA owns
template <typename T, typename XF, int NSTAGES = SymvCfg<T>::STAGES>
__device__ bool symv_cta(SymvSmem<T, NSTAGES>& sm, const T* __restrict__ S, int nT, int pairIdx,
                         const XF& xf, double* __restrict__ rowpart, double* __restrict__ colpart,
                         unsigned int* counter, int ctas_per_matrix, int part = 0, int nparts = 1) {
  // part / nparts: the tiles J of every tile-row are divided into `nparts` contiguous ranges, one
  // CTA each (more CTAs in flight when a GPU holds few workers).  The row sums are then partial
  // as well: `rowpart` must point at this part's own dp-vector and the gather adds the parts.
  constexpr int STAGES = NSTAGES;
  constexpr uint32_t kTileBytes = kT * kT * sizeof(T);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], kConsumerWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int rowA = pairIdx, rowB = nT - 1 - pairIdx;
  const int nrows = (rowA == rowB) ? 1 : 2;

  if (warp == kConsumerWarps) {
    // ------------------------------------------------------------------ producer warp
    // Per tile the producer has to (a) wait for a free stage and issue the bulk copy and (b)
    // assemble the RHS slice x_J (xf: a few global loads + arithmetic, one L2 round trip) and
    // hand the stage to the consumers.  Done tile by tile, the round trip of (b) is the
    // producer's whole period: ncu on the fp32 instantiation showed it spending half of its time
    // there, almost none waiting for a free stage, and the consumers 26 % of theirs on the
    // `full` barriers (profiles/half_iter_fp64.md).  So the tiles of both tile-rows form one flat
    // sequence that is processed in groups of G: the raw loads of a whole group are issued back
    // to back (one round trip per group), then, tile by tile, the bulk copy of the tile one group
    // AHEAD is issued and the tile of this group is handed off -- the arithmetic on the raw
    // values sits behind the (asm volatile) stage wait, so the loads are in flight while the next
    // copies are issued and the ring stays full.  G <= STAGES - 1: issuing tile t + G needs the
    // stage of tile t + G - STAGES back, which must have been handed off already.
    //
    // History (same-box A/Bs at 128 workers x d=4096 per GPU, ms per iteration fp64 / fp32 tiles):
    //   tile by tile                                   1.3026 / 0.7765
    //   register prefetch of computed slices, 4 ahead  9 % slower (fp64)
    //   one- / two-deep pipelining of COMPUTED slices  neutral (the compiler keeps the arithmetic,
    //                                                  hence the stall, next to the loads)
    //   groups of 3 (fp32) / 2 (fp64), no look-ahead   1.2970 / 0.7374
    using XS = XSplit<XF>;
    using Raw = typename XS::Raw;
    constexpr int G = XGroup<T>::G < STAGES - 1 ? XGroup<T>::G : STAGES - 1;
    const int JbA = (int)(((long long)(rowA + 1) * part) / nparts);
    const int JeA = (int)(((long long)(rowA + 1) * (part + 1)) / nparts);
    const int JbB = (int)(((long long)(rowB + 1) * part) / nparts);
    const int JeB = (int)(((long long)(rowB + 1) * (part + 1)) / nparts);
    const int nA = JeA - JbA;
    const int total = nA + (nrows == 2 ? JeB - JbB : 0);
    const T* baseA = S + ((size_t)rowA * (rowA + 1) / 2) * (kT * kT);
    const T* baseB = S + ((size_t)rowB * (rowB + 1) / 2) * (kT * kT);
    int istage = 0;            // ring position of the next bulk copy
    uint32_t iphase = 0;
    int hstage = 0;            // ring position of the next hand-off
    auto issue = [&](int t) {
      const T* src = t < nA ? baseA + (size_t)(JbA + t) * (kT * kT)
                            : baseB + (size_t)(JbB + t - nA) * (kT * kT);
      mbar_wait(&sm.empty[istage], iphase ^ 1);
      if (lane == 0) bulk_g2s(sm.tile[istage], src, kTileBytes, &sm.full[istage]);
      if (++istage == STAGES) { istage = 0; iphase ^= 1; }
    };
#pragma unroll
    for (int g = 0; g < G; ++g)
      if (g < total) issue(g);
    {
      // x_I slices for the column products of the two tile-rows (read after the first `full`)
      const Raw a0 = XS::load(xf, rowA * kT + lane), a1 = XS::load(xf, rowA * kT + lane + 32);
      const Raw b0 = XS::load(xf, rowB * kT + lane), b1 = XS::load(xf, rowB * kT + lane + 32);
      sm.xrow[0][lane] = XS::fin(xf, a0);
      sm.xrow[0][lane + 32] = XS::fin(xf, a1);
      sm.xrow[1][lane] = XS::fin(xf, b0);
      sm.xrow[1][lane + 32] = XS::fin(xf, b1);
    }
    for (int t0 = 0; t0 < total; t0 += G) {
      Raw raw[G][2];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const int t = min(t0 + g, total - 1);    // clamped, not predicated: the loads must stay
        const int J = t < nA ? JbA + t : JbB + t - nA;   // branch-free to be issued back to back
        raw[g][0] = XS::load(xf, J * kT + lane);
        raw[g][1] = XS::load(xf, J * kT + lane + 32);
      }
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (t0 + G + g < total) issue(t0 + G + g);
        if (t0 + g < total) {
          sm.xs[hstage][lane] = XS::fin(xf, raw[g][0]);
          sm.xs[hstage][lane + 32] = XS::fin(xf, raw[g][1]);
          __syncwarp();
          if (lane == 0) mbar_arrive_expect_tx(&sm.full[hstage], kTileBytes);
          if (++hstage == STAGES) hstage = 0;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ consumer warps
    int stage = 0;
    uint32_t phase = 0;
    int t = 0;  // running tile counter (colbuf double buffer)
    for (int rr = 0; rr < nrows; ++rr) {
      const int I = rr == 0 ? rowA : rowB;
      double acc_row[kRowsPerWarp];
      double xr[kRowsPerWarp];
#pragma unroll
      for (int r = 0; r < kRowsPerWarp; ++r) acc_row[r] = 0.0;
      const int Jb = (int)(((long long)(I + 1) * part) / nparts);
      const int Je = (int)(((long long)(I + 1) * (part + 1)) / nparts);
      for (int J = Jb; J < Je; ++J, ++t) {
        mbar_wait(&sm.full[stage], phase);
        if (J == Jb) {
#pragma unroll
          for (int r = 0; r < kRowsPerWarp; ++r) xr[r] = sm.xrow[rr][warp * kRowsPerWarp + r];
        }
        const double2 xj = *reinterpret_cast<const double2*>(&sm.xs[stage][2 * lane]);
        const T* tp = &sm.tile[stage][(warp * kRowsPerWarp) * kT + 2 * lane];
        double2 acc_col = make_double2(0.0, 0.0);
#pragma unroll
        for (int r = 0; r < kRowsPerWarp; ++r) {
          const double2 a = ld_pair(tp + r * kT);
          acc_row[r] = fma(a.x, xj.x, acc_row[r]);
          acc_row[r] = fma(a.y, xj.y, acc_row[r]);
          acc_col.x = fma(a.x, xr[r], acc_col.x);
          acc_col.y = fma(a.y, xr[r], acc_col.y);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        if (J < I) {
          double* cb = sm.colbuf[t & 1][warp];
          *reinterpret_cast<double2*>(&cb[2 * lane]) = acc_col;
          asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
          if (tid < kT) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kConsumerWarps; ++w) s += sm.colbuf[t & 1][w][tid];
            colpart[colpart_off(I, J) + tid] = s;
          }
        }
      }
      // finish the row sums of this tile-row inside the warp
#pragma unroll
      for (int r = 0; r < kRowsPerWarp; ++r) acc_row[r] = warp_sum(acc_row[r]);
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < kRowsPerWarp; ++r)
          rowpart[I * kT + warp * kRowsPerWarp + r] = acc_row[r];
      }
    }
  }
  // ---- ticket: the last CTA of this matrix becomes the finisher
  __threadfence();
  __syncthreads();
  if (tid == 0) sm.ticket = atomicAdd(counter, 1u);
  __syncthreads();
  const bool last = (sm.ticket == (unsigned)(ctas_per_matrix - 1));
  if (last) {
    __threadfence();
    if (tid == 0) *counter = 0u;   // re-arm for the next launch
  }
  return last;
}

// y_j = rowpart[j] + sum_{I > j/64} colpart(I, j/64)[j%64], fixed order (deterministic).
__device__ __forceinline__ double symv_gather(const double* rowpart, const double* colpart, int nT,
                                              int j, int nparts = 1, size_t part_stride = 0) {
  const int J = j / kT, jj = j % kT;
  double v = __ldcg(rowpart + j);
  for (int q = 1; q < nparts; ++q) v += __ldcg(rowpart + q * part_stride + j);
#pragma unroll 8
  for (int I = J + 1; I < nT; ++I) v += __ldcg(colpart + colpart_off(I, J) + jj);   // loads overlap
  return v;
}

}  // namespace gadmm
