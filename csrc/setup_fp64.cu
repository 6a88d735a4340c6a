// One-off setup in fp64 (K1 exact path, K2, pack); block products on the fp64 tensor path (DMMA):
//   Gram  G_n = X_n' X_n, b_n = X_n' y_n, 0.5 y_n' y_n        (group_ADMM_closedForm.m:43-45 `H'*H`)
//   A_n   = G_n + c rho I                                     (same lines, the `\` operand)
//   A_n^-1 by blocked Cholesky (64-wide panels): potrf + explicit inverse of the diagonal block,
//          left-looking column update, panel = A_ik Linv_kk', then W = L^-1 by block forward substitution and
//          A^-1 = W'W.  Everything except the 64x64 diagonal work is a 64x64x64 block product.
//   pack  lower block-triangle -> 64x64 tiles (fp64 / fp32) for the streaming SYMV.
// The reference rebuilds and refactors H'H + c rho I in every iteration of every worker; here it
// happens once per (worker, c).  tcgen05 split-tf32 Gram for the scaled configs: gram_tcgen05.cu.
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

constexpr int B = 64;         // block edge
constexpr int LDS = B + 4;    // padded smem leading dimension (doubles): 68 = 4 mod 16 makes both
                              // fragment access patterns below bank-conflict free
constexpr int kBlkSmem = 2 * B * LDS * (int)sizeof(double);        // Sa + Sb (dynamic)
constexpr int kDiagSmem = 2 * B * (B + 1) * (int)sizeof(double);   // L + Li  (dynamic)

// 64x64 block product on the fp64 tensor path: D(8x8) += A(8x4) B(4x8) per warp instruction
// (mma.sync m8n8k4 f64 -- tcgen05 has no fp64 kind, this is the only tensor-core route for
// doubles).  Versus a 4x4-per-thread FMA tile it needs a quarter of the shared-memory reads per
// multiply-add, which is what bounded the SIMT version.
// 8 warps: warp w owns rows (w>>1)*16..+16 and columns (w&1)*32..+32 of the block, i.e. 2 x 4
// m8n8 tiles; thread (gid = lane>>2, tig = lane&3) holds c[mt][nt][e] = C(row, col) with
//   row = (w>>1)*16 + mt*8 + gid,   col = (w&1)*32 + nt*8 + 2*tig + e.
struct Acc {
  double c[2][4][2];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) c[a][b][0] = c[a][b][1] = 0.0;
  }
};

__device__ __forceinline__ void dmma_m8n8k4(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}

// acc += A * B over klen (multiple of 4) with both operands in shared memory as plain copies of
// their global blocks:  A(m,kk) = A_ROWS_M ? Sa[m][kk] : Sa[kk][m],
//                        B(kk,n) = B_ROWS_N ? Sb[n][kk] : Sb[kk][n].
template <bool A_ROWS_M, bool B_ROWS_N>
__device__ __forceinline__ void block_dmma(const double (*Sa)[LDS], const double (*Sb)[LDS], Acc& acc,
                                           int klen = B) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tig = lane & 3;
  const int wr = (warp >> 1) * 16, wc = (warp & 1) * 32;
#pragma unroll 4
  for (int k0 = 0; k0 < klen; k0 += 4) {
    double a[2], b[4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
      a[mt] = A_ROWS_M ? Sa[wr + mt * 8 + gid][k0 + tig] : Sa[k0 + tig][wr + mt * 8 + gid];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
      b[nt] = B_ROWS_N ? Sb[wc + nt * 8 + gid][k0 + tig] : Sb[k0 + tig][wc + nt * 8 + gid];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) dmma_m8n8k4(acc.c[mt][nt], a[mt], b[nt]);
  }
}

// f(row, col, value&) over the thread's 16 accumulator entries (block-local coordinates)
template <typename F>
__device__ __forceinline__ void acc_for_each(Acc& acc, F&& f) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tig = lane & 3;
  const int wr = (warp >> 1) * 16, wc = (warp & 1) * 32;
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) f(wr + mt * 8 + gid, wc + nt * 8 + 2 * tig + e, acc.c[mt][nt][e]);
}

// Copy a 64x64 global block into shared memory as is: dst[r][c] = p[r*ld + c] (coalesced,
// conflict-free), optionally scaled.
__device__ __forceinline__ void load_block(double (*dst)[LDS], const double* __restrict__ p,
                                           size_t ld, double scale = 1.0) {
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int idx = t + i * 256, r = idx >> 6, c = idx & 63;
    dst[r][c] = scale * p[(size_t)r * ld + c];
  }
}

// ------------------------------------------------------------------------------------- Gram
// grid: (lower-triangular tile pairs, N); X_n is s x d row-major.  G tile (I,J) and its mirror.
__global__ void __launch_bounds__(256)
gram_f64_kernel(const double* __restrict__ X, int s, int d, int dp, double* __restrict__ G) {
  __shared__ double As[16][LDS], Bs[16][LDS];
  const int n = blockIdx.y;
  int I = 0, rem = blockIdx.x;
  while (rem > I) { rem -= I + 1; ++I; }
  const int J = rem;
  const double* Xn = X + (size_t)n * s * d;
  const int t = threadIdx.x;
  Acc acc;
  acc.zero();
  for (int k0 = 0; k0 < s; k0 += 16) {
    // 16 x 64 slices of X for the I columns and the J columns (coalesced along features)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = t + i * 256, kk = idx >> 6, m = idx & 63;
      const int row = k0 + kk, ci = I * B + m, cj = J * B + m;
      As[kk][m] = (row < s && ci < d) ? Xn[(size_t)row * d + ci] : 0.0;
      Bs[kk][m] = (row < s && cj < d) ? Xn[(size_t)row * d + cj] : 0.0;
    }
    __syncthreads();
    block_dmma<false, false>(As, Bs, acc, 16);      // G_IJ(m,nn) += X(kk, I m) X(kk, J nn)
    __syncthreads();
  }
  double* Gn = G + (size_t)n * dp * dp;
  acc_for_each(acc, [&](int r, int c, double& v) {
    const int gr = I * B + r, gc = J * B + c;
    Gn[(size_t)gr * dp + gc] = v;
    Gn[(size_t)gc * dp + gr] = v;
  });
}

// b_n = X_n' y_n (one CTA per (n, 64-feature block)), hyy_n = 0.5 y'y
__global__ void __launch_bounds__(256)
xty_kernel(const double* __restrict__ X, const double* __restrict__ y, int s, int d, int dp,
           double* __restrict__ b, double* __restrict__ hyy) {
  __shared__ double part[4][64];
  __shared__ double red[32];
  const int n = blockIdx.y, c0 = blockIdx.x * 64;
  const double* Xn = X + (size_t)n * s * d;
  const double* yn = y + (size_t)n * s;
  const int col = threadIdx.x & 63, grp = threadIdx.x >> 6;
  double acc = 0.0;
  if (c0 + col < d)
    for (int r = grp; r < s; r += 4) acc = fma(Xn[(size_t)r * d + c0 + col], yn[r], acc);
  part[grp][col] = acc;
  __syncthreads();
  if (threadIdx.x < 64)
    b[(size_t)n * dp + c0 + col] = (part[0][col] + part[1][col]) + (part[2][col] + part[3][col]);
  if (blockIdx.x == 0) {
    double v = 0.0;
    for (int r = threadIdx.x; r < s; r += blockDim.x) v = fma(yn[r], yn[r], v);
    v = block_sum(v, red);
    if (threadIdx.x == 0) hyy[n] = 0.5 * v;
  }
}

__global__ void shift_copy_kernel(const double* __restrict__ G, double* __restrict__ A, int d,
                                  int dp, double shift) {
  const size_t n = blockIdx.y;
  const size_t tot = (size_t)dp * dp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < tot;
       i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / dp), c = (int)(i % dp);
    double v = (r < d && c < d) ? G[n * tot + i] : 0.0;
    if (r == c) v = (r < d) ? v + shift : 1.0;
    A[n * tot + i] = v;
  }
}

// ------------------------------------------------------------------------ blocked Cholesky
// Diagonal block k: L_kk = chol(A_kk) in place (upper zeroed) and Dinv[n][k] = L_kk^-1.
// 256 threads: t = row, q = one of four column groups of the trailing update (this kernel sits on
// the critical path of every block column, so its latency matters more than its flops).
__global__ void __launch_bounds__(256)
potrf_diag_kernel(double* __restrict__ A, double* __restrict__ Dinv, int dp, int nT, int k) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*L)[B + 1] = reinterpret_cast<double (*)[B + 1]>(dsm);
  double (*Li)[B + 1] = L + B;
  const int n = blockIdx.x, t = threadIdx.x & 63, q = threadIdx.x >> 6;
  double* Ak = A + (size_t)n * dp * dp + (size_t)k * B * dp + (size_t)k * B;
  for (int r = q; r < B; r += 4) {
    L[r][t] = Ak[(size_t)r * dp + t];
    Li[r][t] = 0.0;
  }
  __syncthreads();
  for (int j = 0; j < B; ++j) {
    if (threadIdx.x == 0) L[j][j] = sqrt(L[j][j]);
    __syncthreads();
    if (q == 0 && t > j) L[t][j] /= L[j][j];
    __syncthreads();
    // rank-1 update of the trailing lower part: row t, columns j+1..t, split over the 4 groups
    if (t > j) {
      const double ltj = L[t][j];
      for (int cc = j + 1 + q; cc <= t; cc += 4) L[t][cc] = fma(-ltj, L[cc][j], L[t][cc]);
    }
    __syncthreads();
  }
  // inverse of the lower-triangular block: thread t (group 0) solves L x = e_t
  if (q == 0) {
    for (int r = t; r < B; ++r) {
      double v = (r == t) ? 1.0 : 0.0;
      for (int m = t; m < r; ++m) v = fma(-L[r][m], Li[m][t], v);
      Li[r][t] = v / L[r][r];
    }
  }
  __syncthreads();
  double* Dk = Dinv + ((size_t)n * nT + k) * B * B;
  for (int r = q; r < B; r += 4) {
    Ak[(size_t)r * dp + t] = (t <= r) ? L[r][t] : 0.0;
    Dk[r * B + t] = Li[r][t];
  }
}

// Left-looking block column j, tiles i >= j:  A_ij -= sum_{k<j} L_ik L_jk'  accumulated in
// registers over the whole k range (one read-modify-write of A_ij per factorisation instead of one
// per panel, which made the right-looking version memory bound).  grid: (nT - j, N)
__global__ void __launch_bounds__(256)
chol_update_kernel(double* __restrict__ A, int dp, int nT, int j) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*Sa)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Sb)[LDS] = Sa + B;
  const int n = blockIdx.y, i = j + blockIdx.x;
  double* An = A + (size_t)n * dp * dp;
  Acc acc;
  acc.zero();
  for (int k = 0; k < j; ++k) {
    load_block(Sa, An + (size_t)i * B * dp + (size_t)k * B, dp);   // L_ik(m,kk)
    load_block(Sb, An + (size_t)j * B * dp + (size_t)k * B, dp);   // L_jk(nn,kk)
    __syncthreads();
    block_dmma<true, true>(Sa, Sb, acc);
    __syncthreads();
  }
  double* Cij = An + (size_t)i * B * dp + (size_t)j * B;
  acc_for_each(acc, [&](int r, int c, double& v) { Cij[(size_t)r * dp + c] -= v; });
}

// Panel: L_ij = A_ij * Linv_jj'  for i = j+1..nT-1   (grid: (nT-j-1, N))
__global__ void __launch_bounds__(256)
panel_kernel(double* __restrict__ A, const double* __restrict__ Dinv, int dp, int nT, int k) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*Sa)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Sb)[LDS] = Sa + B;
  const int n = blockIdx.y, i = k + 1 + blockIdx.x;
  double* Aik = A + (size_t)n * dp * dp + (size_t)i * B * dp + (size_t)k * B;
  const double* Dk = Dinv + ((size_t)n * nT + k) * B * B;
  // C(m,nn) = sum_kk A_ik(m,kk) * Linv(nn,kk)
  load_block(Sa, Aik, dp);
  load_block(Sb, Dk, B);
  __syncthreads();
  Acc acc;
  acc.zero();
  block_dmma<true, true>(Sa, Sb, acc);
  acc_for_each(acc, [&](int r, int c, double& v) { Aik[(size_t)r * dp + c] = v; });
}

// Row i of W = L^-1:  W_ik = -Linv_ii * sum_{j=k}^{i-1} L_ij W_jk  (k < i),  W_ii = Linv_ii.
// grid: (i, N)  -- blockIdx.x = k in [0, i)
__global__ void __launch_bounds__(256)
trtri_row_kernel(const double* __restrict__ L, double* __restrict__ W,
                 const double* __restrict__ Dinv, int dp, int nT, int i) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*Sa)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Sb)[LDS] = Sa + B;
  const int n = blockIdx.y, k = blockIdx.x;
  const double* Ln = L + (size_t)n * dp * dp;
  double* Wn = W + (size_t)n * dp * dp;
  Acc acc;
  acc.zero();
  for (int j = k; j < i; ++j) {
    load_block(Sa, Ln + (size_t)i * B * dp + (size_t)j * B, dp);   // L_ij(m,kk)
    load_block(Sb, Wn + (size_t)j * B * dp + (size_t)k * B, dp);   // W_jk(kk,nn)
    __syncthreads();
    block_dmma<true, false>(Sa, Sb, acc);
    __syncthreads();
  }
  // T = acc (64x64) -> Sb[kk][nn], then W_ik = -Linv_ii * T
  acc_for_each(acc, [&](int r, int c, double& v) { Sb[r][c] = v; });
  load_block(Sa, Dinv + ((size_t)n * nT + i) * B * B, B, -1.0);    // -Linv_ii(m,kk)
  __syncthreads();
  acc.zero();
  block_dmma<true, false>(Sa, Sb, acc);
  double* Wik = Wn + (size_t)i * B * dp + (size_t)k * B;
  acc_for_each(acc, [&](int r, int c, double& v) { Wik[(size_t)r * dp + c] = v; });
}

// W diagonal blocks = Dinv (grid: (nT, N), 256 threads)
__global__ void __launch_bounds__(256)
trtri_diag_kernel(double* __restrict__ W, const double* __restrict__ Dinv, int dp, int nT) {
  const int n = blockIdx.y, k = blockIdx.x;
  const double* Dk = Dinv + ((size_t)n * nT + k) * B * B;
  double* Wkk = W + (size_t)n * dp * dp + (size_t)k * B * dp + (size_t)k * B;
  for (int idx = threadIdx.x; idx < B * B; idx += 256) Wkk[(size_t)(idx >> 6) * dp + (idx & 63)] = Dk[idx];
}

// M_ij = sum_{k>=i} W_ki' W_kj   for i >= j   (grid: (pairs, N)); written into A (L is dead)
__global__ void __launch_bounds__(256)
lauum_kernel(const double* __restrict__ W, double* __restrict__ M, int dp, int nT) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*Sa)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Sb)[LDS] = Sa + B;
  const int n = blockIdx.y;
  int i = 0, rem = blockIdx.x;
  while (rem > i) { rem -= i + 1; ++i; }
  const int j = rem;
  const double* Wn = W + (size_t)n * dp * dp;
  Acc acc;
  acc.zero();
  for (int k = i; k < nT; ++k) {
    load_block(Sa, Wn + (size_t)k * B * dp + (size_t)i * B, dp);   // W_ki(kk,m)
    load_block(Sb, Wn + (size_t)k * B * dp + (size_t)j * B, dp);   // W_kj(kk,nn)
    __syncthreads();
    block_dmma<false, false>(Sa, Sb, acc);
    __syncthreads();
  }
  double* Mij = M + (size_t)n * dp * dp + (size_t)i * B * dp + (size_t)j * B;
  acc_for_each(acc, [&](int r, int c, double& v) { Mij[(size_t)r * dp + c] = v; });
}

// ------------------------------------------------------------------------------------- pack
template <typename T>
__global__ void __launch_bounds__(256)
pack_sym_kernel(const double* __restrict__ M, T* __restrict__ P, int d, int dp, size_t packed) {
  const int n = blockIdx.y;
  int I = 0, rem = blockIdx.x;
  while (rem > I) { rem -= I + 1; ++I; }
  const int J = rem;
  const double* Mn = M + (size_t)n * dp * dp;
  T* tile = P + (size_t)n * packed + (size_t)blockIdx.x * (B * B);
  for (int idx = threadIdx.x; idx < B * B; idx += 256) {
    const int r = idx >> 6, c = idx & 63;
    int gr = I * B + r, gc = J * B + c;
    if (I == J && c > r) { const int tmp = gr; gr = gc; gc = tmp; }   // mirror: only lower is valid
    const double v = (gr < d && gc < d) ? Mn[(size_t)gr * dp + gc] : 0.0;
    tile[idx] = (T)v;
  }
}

}  // namespace gadmm

using namespace gadmm;

extern "C" {

int gadmm_gram_f64(const double* X, const double* y, int N, int s, int d, int dp, double* G,
                   double* b, double* hyy, void* stream) {
  GADMM_REQUIRE(dp % B == 0 && dp >= d && d > 0 && s > 0, "gram: bad shape s=%d d=%d dp=%d", s, d, dp);
  if (N <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int nT = dp / B;
  GADMM_CUDA_OK(cudaMemsetAsync(b, 0, (size_t)N * dp * sizeof(double), st));
  gram_f64_kernel<<<dim3(nT * (nT + 1) / 2, N), 256, 0, st>>>(X, s, d, dp, G);
  xty_kernel<<<dim3(nT, N), 256, 0, st>>>(X, y, s, d, dp, b, hyy);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_shift_copy(const double* G, double* A, int N, int d, int dp, double shift, void* stream) {
  if (N <= 0) return 0;
  shift_copy_kernel<<<dim3(64, N), 256, 0, (cudaStream_t)stream>>>(G, A, d, dp, shift);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_spd_inverse_f64(double* A, double* W, double* Dinv, int N, int dp, void* stream) {
  GADMM_REQUIRE(dp > 0 && dp % B == 0, "spd_inverse: dp=%d must be a multiple of 64", dp);
  if (N <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int nT = dp / B;
  GADMM_CUDA_OK(cudaFuncSetAttribute(potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDiagSmem));
  GADMM_CUDA_OK(cudaFuncSetAttribute(panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  GADMM_CUDA_OK(cudaFuncSetAttribute(chol_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  GADMM_CUDA_OK(cudaFuncSetAttribute(trtri_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  GADMM_CUDA_OK(cudaFuncSetAttribute(lauum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  for (int k = 0; k < nT; ++k) {       // left-looking: bring column k up to date, factor, solve
    if (k > 0) chol_update_kernel<<<dim3(nT - k, N), 256, kBlkSmem, st>>>(A, dp, nT, k);
    potrf_diag_kernel<<<N, 256, kDiagSmem, st>>>(A, Dinv, dp, nT, k);
    const int rest = nT - k - 1;
    if (rest > 0) panel_kernel<<<dim3(rest, N), 256, kBlkSmem, st>>>(A, Dinv, dp, nT, k);
  }
  GADMM_CUDA_OK(cudaMemsetAsync(W, 0, (size_t)N * dp * dp * sizeof(double), st));
  trtri_diag_kernel<<<dim3(nT, N), 256, 0, st>>>(W, Dinv, dp, nT);
  for (int i = 1; i < nT; ++i)
    trtri_row_kernel<<<dim3(i, N), 256, kBlkSmem, st>>>(A, W, Dinv, dp, nT, i);
  lauum_kernel<<<dim3(nT * (nT + 1) / 2, N), 256, kBlkSmem, st>>>(W, A, dp, nT);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int gadmm_pack_sym(const double* M, void* packed, int N, int d, int dp, int store, void* stream) {
  GADMM_REQUIRE(dp > 0 && dp % B == 0, "pack: dp=%d must be a multiple of 64", dp);
  if (N <= 0) return 0;
  const int nT = dp / B;
  const size_t pe = gadmm_packed_elems(dp);
  dim3 grid(nT * (nT + 1) / 2, N);
  if (store == GADMM_STORE_F64)
    pack_sym_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(M, (double*)packed, d, dp, pe);
  else if (store == GADMM_STORE_F32)
    pack_sym_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(M, (float*)packed, d, dp, pe);
  else
    return set_error(-1, "pack: bad store %d", store);
  GADMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
