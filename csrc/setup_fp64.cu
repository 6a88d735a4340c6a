// One-off setup in fp64 SIMT (K1 exact path, K2, pack):
//   Gram  G_n = X_n' X_n, b_n = X_n' y_n, 0.5 y_n' y_n        (group_ADMM_closedForm.m:43-45 `H'*H`)
//   A_n   = G_n + c rho I                                     (same lines, the `\` operand)
//   A_n^-1 by blocked Cholesky (64-wide panels): potrf + explicit inverse of the diagonal block,
//          panel = A_ik Linv_kk', trailing SYRK, then W = L^-1 by block forward substitution and
//          A^-1 = W'W.  Everything except the 64x64 diagonal work is a 64x64x64 block product.
//   pack  lower block-triangle -> 64x64 tiles (fp64 / fp32) for the streaming SYMV.
// The reference rebuilds and refactors H'H + c rho I in every iteration of every worker; here it
// happens once per (worker, c).  tcgen05 split-tf32 Gram for the scaled configs: gram_tcgen05.cu.
#include "../include/gadmm_b200.h"
#include "common.cuh"

namespace gadmm {

constexpr int B = 64;         // block edge
constexpr int LDS = B + 2;    // padded smem leading dimension (doubles)
constexpr int kBlkSmem = 2 * B * LDS * (int)sizeof(double);        // As + Bs (dynamic)
constexpr int kDiagSmem = 2 * B * (B + 1) * (int)sizeof(double);   // L + Li  (dynamic)

// acc(4x4 per thread, 256 threads) += As' * Bs  where As[kk][m], Bs[kk][n] (k-major in smem)
__device__ __forceinline__ void block_mma(const double (*As)[LDS], const double (*Bs)[LDS],
                                          double acc[4][4], int ty, int tx, int klen = B) {
#pragma unroll 8
  for (int kk = 0; kk < klen; ++kk) {
    double a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
  }
}

// Load a 64x64 block into smem k-major: dst[kk][m] = src(m, kk).
//   trans == false: src(m,kk) = p[m*ld + kk]   (row-major block, k contiguous)
//   trans == true : src(m,kk) = p[kk*ld + m]   (m contiguous)
__device__ __forceinline__ void load_block(double (*dst)[LDS], const double* __restrict__ p,
                                           size_t ld, bool trans, double scale = 1.0) {
  const int t = threadIdx.x;
  if (trans) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = t + i * 256, kk = idx >> 6, m = idx & 63;
      dst[kk][m] = scale * p[(size_t)kk * ld + m];
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = t + i * 256, m = idx >> 6, kk = idx & 63;
      dst[kk][m] = scale * p[(size_t)m * ld + kk];
    }
  }
}

__device__ __forceinline__ void zero_acc(double acc[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
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
  const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
  double acc[4][4];
  zero_acc(acc);
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
    block_mma(As, Bs, acc, ty, tx, 16);
    __syncthreads();
  }
  double* Gn = G + (size_t)n * dp * dp;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = I * B + ty * 4 + i, cidx = J * B + tx * 4 + j;
      Gn[(size_t)r * dp + cidx] = acc[i][j];
      Gn[(size_t)cidx * dp + r] = acc[i][j];
    }
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
__global__ void __launch_bounds__(64)
potrf_diag_kernel(double* __restrict__ A, double* __restrict__ Dinv, int dp, int nT, int k) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*L)[B + 1] = reinterpret_cast<double (*)[B + 1]>(dsm);
  double (*Li)[B + 1] = L + B;
  const int n = blockIdx.x, t = threadIdx.x;
  double* Ak = A + (size_t)n * dp * dp + (size_t)k * B * dp + (size_t)k * B;
  for (int r = 0; r < B; ++r) L[r][t] = Ak[(size_t)r * dp + t];
  __syncthreads();
  for (int j = 0; j < B; ++j) {
    if (t == j) L[j][j] = sqrt(L[j][j]);
    __syncthreads();
    if (t > j) L[t][j] /= L[j][j];
    __syncthreads();
    // rank-1 update of the trailing lower part: row t, columns j+1..t
    if (t > j)
      for (int cc = j + 1; cc <= t; ++cc) L[t][cc] = fma(-L[t][j], L[cc][j], L[t][cc]);
    __syncthreads();
  }
  // inverse of the lower-triangular block: thread t solves L x = e_t
  for (int r = 0; r < B; ++r) Li[r][t] = 0.0;
  __syncthreads();
  for (int r = t; r < B; ++r) {
    double v = (r == t) ? 1.0 : 0.0;
    for (int m = t; m < r; ++m) v = fma(-L[r][m], Li[m][t], v);
    Li[r][t] = v / L[r][r];
  }
  __syncthreads();
  double* Dk = Dinv + ((size_t)n * nT + k) * B * B;
  for (int r = 0; r < B; ++r) {
    Ak[(size_t)r * dp + t] = (t <= r) ? L[r][t] : 0.0;
    Dk[r * B + t] = Li[r][t];
  }
}

// Panel: L_ik = A_ik * Linv_kk'  for i = k+1..nT-1   (grid: (nT-k-1, N))
__global__ void __launch_bounds__(256)
panel_kernel(double* __restrict__ A, const double* __restrict__ Dinv, int dp, int nT, int k) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*As)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Bs)[LDS] = As + B;
  const int n = blockIdx.y, i = k + 1 + blockIdx.x;
  double* Aik = A + (size_t)n * dp * dp + (size_t)i * B * dp + (size_t)k * B;
  const double* Dk = Dinv + ((size_t)n * nT + k) * B * B;
  // C(m,nn) = sum_kk A_ik(m,kk) * Linv(nn,kk)
  load_block(As, Aik, dp, false);
  load_block(Bs, Dk, B, false);     // Bs[kk][nn] = Linv(nn,kk)
  __syncthreads();
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  double acc[4][4];
  zero_acc(acc);
  block_mma(As, Bs, acc, ty, tx);
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b2 = 0; b2 < 4; ++b2) Aik[(size_t)(ty * 4 + a) * dp + tx * 4 + b2] = acc[a][b2];
}

// Trailing update: A_ij -= L_ik L_jk'   for k < j <= i   (grid: (pairs, N))
__global__ void __launch_bounds__(256)
syrk_kernel(double* __restrict__ A, int dp, int nT, int k) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*As)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Bs)[LDS] = As + B;
  const int n = blockIdx.y;
  int ii = 0, rem = blockIdx.x;
  while (rem > ii) { rem -= ii + 1; ++ii; }
  const int i = k + 1 + ii, j = k + 1 + rem;
  double* An = A + (size_t)n * dp * dp;
  load_block(As, An + (size_t)i * B * dp + (size_t)k * B, dp, false);
  load_block(Bs, An + (size_t)j * B * dp + (size_t)k * B, dp, false);
  __syncthreads();
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  double acc[4][4];
  zero_acc(acc);
  block_mma(As, Bs, acc, ty, tx);
  double* Cij = An + (size_t)i * B * dp + (size_t)j * B;
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b2 = 0; b2 < 4; ++b2) Cij[(size_t)(ty * 4 + a) * dp + tx * 4 + b2] -= acc[a][b2];
}

// Row i of W = L^-1:  W_ik = -Linv_ii * sum_{j=k}^{i-1} L_ij W_jk  (k < i),  W_ii = Linv_ii.
// grid: (i, N)  -- blockIdx.x = k in [0, i)
__global__ void __launch_bounds__(256)
trtri_row_kernel(const double* __restrict__ L, double* __restrict__ W,
                 const double* __restrict__ Dinv, int dp, int nT, int i) {
  extern __shared__ __align__(16) unsigned char dsm[];
  double (*As)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Bs)[LDS] = As + B;
  const int n = blockIdx.y, k = blockIdx.x;
  const double* Ln = L + (size_t)n * dp * dp;
  double* Wn = W + (size_t)n * dp * dp;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  double acc[4][4];
  zero_acc(acc);
  for (int j = k; j < i; ++j) {
    load_block(As, Ln + (size_t)i * B * dp + (size_t)j * B, dp, false);   // L_ij(m,kk)
    load_block(Bs, Wn + (size_t)j * B * dp + (size_t)k * B, dp, true);    // W_jk(kk,nn)
    __syncthreads();
    block_mma(As, Bs, acc, ty, tx);
    __syncthreads();
  }
  // T = acc (64x64) -> smem (k-major: Bs[kk][nn] = T(kk,nn)), then W_ik = -Linv_ii * T
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b2 = 0; b2 < 4; ++b2) Bs[ty * 4 + a][tx * 4 + b2] = acc[a][b2];
  load_block(As, Dinv + ((size_t)n * nT + i) * B * B, B, false, -1.0);   // As[kk][m] = -Linv(m,kk)
  __syncthreads();
  zero_acc(acc);
  block_mma(As, Bs, acc, ty, tx);
  double* Wik = Wn + (size_t)i * B * dp + (size_t)k * B;
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b2 = 0; b2 < 4; ++b2) Wik[(size_t)(ty * 4 + a) * dp + tx * 4 + b2] = acc[a][b2];
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
  double (*As)[LDS] = reinterpret_cast<double (*)[LDS]>(dsm);
  double (*Bs)[LDS] = As + B;
  const int n = blockIdx.y;
  int i = 0, rem = blockIdx.x;
  while (rem > i) { rem -= i + 1; ++i; }
  const int j = rem;
  const double* Wn = W + (size_t)n * dp * dp;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  double acc[4][4];
  zero_acc(acc);
  for (int k = i; k < nT; ++k) {
    load_block(As, Wn + (size_t)k * B * dp + (size_t)i * B, dp, true);   // As[kk][m] = W_ki(kk,m)
    load_block(Bs, Wn + (size_t)k * B * dp + (size_t)j * B, dp, true);   // Bs[kk][nn] = W_kj(kk,nn)
    __syncthreads();
    block_mma(As, Bs, acc, ty, tx);
    __syncthreads();
  }
  double* Mij = M + (size_t)n * dp * dp + (size_t)i * B * dp + (size_t)j * B;
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b2 = 0; b2 < 4; ++b2) Mij[(size_t)(ty * 4 + a) * dp + tx * 4 + b2] = acc[a][b2];
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
  GADMM_CUDA_OK(cudaFuncSetAttribute(syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  GADMM_CUDA_OK(cudaFuncSetAttribute(trtri_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  GADMM_CUDA_OK(cudaFuncSetAttribute(lauum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBlkSmem));
  for (int k = 0; k < nT; ++k) {
    potrf_diag_kernel<<<N, 64, kDiagSmem, st>>>(A, Dinv, dp, nT, k);
    const int rest = nT - k - 1;
    if (rest > 0) {
      panel_kernel<<<dim3(rest, N), 256, kBlkSmem, st>>>(A, Dinv, dp, nT, k);
      syrk_kernel<<<dim3(rest * (rest + 1) / 2, N), 256, kBlkSmem, st>>>(A, dp, nT, k);
    }
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
