// Packed-symmetric streaming SYMV core (one CTA = two tile-rows of one matrix).
//
// Layout: lower block-triangle of 64x64 row-major tiles, tile (I,J) at index I(I+1)/2+J, diagonal
// tiles stored full.  Reading only the lower triangle halves the HBM bytes of y = S x, which is
// the whole cost of a GADMM half-iteration (d^2 bytes per worker, memory bound).
//
// CTA structure (288 threads): warp 8 is the producer -- it streams tiles global -> shared with
// cp.async.bulk (UBLKCP) through a STAGES-deep mbarrier ring and assembles the matching 64-entry
// slice of x (the fused RHS) beside every tile; warps 0..7 are consumers -- warp w owns rows
// 8w..8w+7 of every tile, lane l owns columns 2l,2l+1.  Per element: one FMA into the row sum
// (y_I += T x_J) and one into the column sum (y_J += T' x_I), all in fp64, no shuffles in the
// inner loop.  Row sums are finished inside the CTA; column sums go to `colpart` and are added
// by the finisher (the last CTA of the matrix, found with an atomic ticket) in a fixed order, so
// results are bit-reproducible run to run.
#pragma once
#include "common.cuh"

namespace gadmm {

constexpr int kT = GADMM_TILE;           // 64
constexpr int kConsumerWarps = 8;
constexpr int kSymvThreads = (kConsumerWarps + 1) * 32;  // 288
constexpr int kRowsPerWarp = kT / kConsumerWarps;        // 8
#ifndef GADMM_BULK_SPLIT
#define GADMM_BULK_SPLIT 1      // splitting a tile over several issuing lanes made no difference
#endif
constexpr int kBulkSplit = GADMM_BULK_SPLIT;

template <typename T> struct SymvCfg;
template <> struct SymvCfg<double> { static constexpr int STAGES = 3; };
template <> struct SymvCfg<float>  { static constexpr int STAGES = 6; };

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
// fp32 -> fp64 widening.  The F2F.F64.F32 instruction runs on the quarter-rate XU pipe, which the
// fp32-tile kernel loads to 34 % on average with bursts above its peak (ncu,
// profiles/half_iter_f32_ncu_raw_r2.csv); GADMM_F2D_INT=1 does the widening with integer ops on
// the ALU pipe instead (exact for normal numbers and zero; fp32 denormals flush to zero, which for
// entries of an inverse matrix is below any effect).
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
    // The x slices are fetched kXG tiles ahead in registers (group g+1 is in flight while the
    // tiles of group g are issued), so the global-load latency of the fused RHS assembly never
    // sits between two bulk copies -- that matters when a GPU has few CTAs in flight (one worker
    // per GPU: 128 CTAs, each needing ~50 GB/s).
// Register prefetch of the x slices one group ahead: measured 9 % SLOWER on the flagship shape
// (1.464 vs 1.331 ms/iteration at 128 workers x d=4096 per GPU; same box A/B), so it is off.
#ifndef GADMM_XPREFETCH
#define GADMM_XPREFETCH 0
#endif
// One-deep software pipeline of the RHS slice loads (GADMM_XPIPE=1): same-box A/B at 128 workers x
// d=4096 per GPU: fp64 tiles 1.3099 vs 1.3096 ms, fp32 tiles 0.785 vs 0.775 ms per iteration --
// the producer is not the limiter, so it is off.  (A two-deep variant measured 1.304 / 0.780 ms
// against 1.300 / 0.766 ms on another box: same conclusion.)
#ifndef GADMM_XPIPE
#define GADMM_XPIPE 0
#endif
    constexpr int kXG = 4;
    int stage = 0;
    uint32_t phase = 0;
#if GADMM_XPIPE
    // One-deep software pipeline over the tiles of BOTH tile-rows: the bulk copy of tile t and the
    // global loads of its RHS slice are issued first, then the slice of tile t-1 (whose loads
    // have had a whole tile period to land) is written to shared memory and tile t-1 is handed
    // to the consumers.  Without it every tile costs one L2 round trip of the fused RHS assembly
    // between two bulk copies (~0.8 us): harmless with 32 KiB fp64 tiles, but it caps a CTA at
    // ~20 GB/s with 16 KiB fp32 tiles.
    {
      double pend0 = 0.0, pend1 = 0.0;
      int pend_stage = -1;
      for (int rr = 0; rr < nrows; ++rr) {
        const int I = rr == 0 ? rowA : rowB;
        const int Jb = (int)(((long long)(I + 1) * part) / nparts);
        const int Je = (int)(((long long)(I + 1) * (part + 1)) / nparts);
        sm.xrow[rr][lane] = xf(I * kT + lane);
        sm.xrow[rr][lane + 32] = xf(I * kT + lane + 32);
        const T* rowbase = S + ((size_t)I * (I + 1) / 2) * (kT * kT);
        for (int J = Jb; J < Je; ++J) {
          mbar_wait(&sm.empty[stage], phase ^ 1);
          if (lane == 0) bulk_g2s(sm.tile[stage], rowbase + (size_t)J * (kT * kT), kTileBytes, &sm.full[stage]);
          const double x0 = xf(J * kT + lane), x1 = xf(J * kT + lane + 32);   // in flight
          if (pend_stage >= 0) {
            sm.xs[pend_stage][lane] = pend0;
            sm.xs[pend_stage][lane + 32] = pend1;
            __syncwarp();
            if (lane == 0) mbar_arrive_expect_tx(&sm.full[pend_stage], kTileBytes);
          }
          pend0 = x0;
          pend1 = x1;
          pend_stage = stage;
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
      if (pend_stage >= 0) {
        sm.xs[pend_stage][lane] = pend0;
        sm.xs[pend_stage][lane + 32] = pend1;
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(&sm.full[pend_stage], kTileBytes);
      }
    }
#elif !GADMM_XPREFETCH
    for (int rr = 0; rr < nrows; ++rr) {
      const int I = rr == 0 ? rowA : rowB;
      const int Jb = (int)(((long long)(I + 1) * part) / nparts);
      const int Je = (int)(((long long)(I + 1) * (part + 1)) / nparts);
      sm.xrow[rr][lane] = xf(I * kT + lane);
      sm.xrow[rr][lane + 32] = xf(I * kT + lane + 32);
      const T* rowbase = S + ((size_t)I * (I + 1) / 2) * (kT * kT);
      for (int J = Jb; J < Je; ++J) {
        mbar_wait(&sm.empty[stage], phase ^ 1);
        if (lane == 0) bulk_g2s(sm.tile[stage], rowbase + (size_t)J * (kT * kT), kTileBytes, &sm.full[stage]);
        sm.xs[stage][lane] = xf(J * kT + lane);
        sm.xs[stage][lane + 32] = xf(J * kT + lane + 32);
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(&sm.full[stage], kTileBytes);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
#else
    for (int rr = 0; rr < nrows; ++rr) {
      const int I = rr == 0 ? rowA : rowB;
      const int Jb = (int)(((long long)(I + 1) * part) / nparts);
      const int Je = (int)(((long long)(I + 1) * (part + 1)) / nparts);     // tiles [Jb, Je)
      double cur[2 * kXG], nxt[2 * kXG];
#pragma unroll
      for (int q = 0; q < kXG; ++q) {
        const bool ok = Jb + q < Je;
        nxt[2 * q] = ok ? xf((Jb + q) * kT + lane) : 0.0;
        nxt[2 * q + 1] = ok ? xf((Jb + q) * kT + lane + 32) : 0.0;
      }
      // x_I slice for the column products of this tile-row
      sm.xrow[rr][lane] = xf(I * kT + lane);
      sm.xrow[rr][lane + 32] = xf(I * kT + lane + 32);
      const T* rowbase = S + ((size_t)I * (I + 1) / 2) * (kT * kT);
      for (int J0 = Jb; J0 < Je; J0 += kXG) {
#pragma unroll
        for (int q = 0; q < 2 * kXG; ++q) cur[q] = nxt[q];
#pragma unroll
        for (int q = 0; q < kXG; ++q) {            // prefetch the next group's slices
          const int Jn = J0 + kXG + q;
          const bool ok = Jn < Je;
          nxt[2 * q] = ok ? xf(Jn * kT + lane) : 0.0;
          nxt[2 * q + 1] = ok ? xf(Jn * kT + lane + 32) : 0.0;
        }
#pragma unroll
        for (int q = 0; q < kXG; ++q) {
          const int J = J0 + q;
          if (J >= Je) break;
          mbar_wait(&sm.empty[stage], phase ^ 1);
          // one tile = kBulkSplit bulk copies issued by different lanes: copies issued by one
          // thread were observed to execute back to back (~25 GB/s per CTA), which starves
          // phases with few CTAs per GPU; independent issuers overlap
          if (lane < kBulkSplit)
            bulk_g2s(reinterpret_cast<unsigned char*>(sm.tile[stage]) + lane * (kTileBytes / kBulkSplit),
                     reinterpret_cast<const unsigned char*>(rowbase + (size_t)J * (kT * kT)) +
                         lane * (kTileBytes / kBulkSplit),
                     kTileBytes / kBulkSplit, &sm.full[stage]);
          sm.xs[stage][lane] = cur[2 * q];
          sm.xs[stage][lane + 32] = cur[2 * q + 1];
          __syncwarp();
          if (lane == 0) mbar_arrive_expect_tx(&sm.full[stage], kTileBytes);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
#endif
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
