/* Minimal C client of the gadmm_b200 C ABI (include/gadmm_b200.h): batched packed-symmetric
 * products y_n = S_n x_n through the allocation-free entry point, on plain cudaMalloc'ed buffers.
 *
 *   gcc -std=c99 -I include -I /usr/local/cuda/include examples/c_api_demo.c \
 *       -L gadmm_b200/_native -lgadmm_b200 -L /usr/local/cuda/lib64 -lcudart \
 *       -Wl,-rpath,$PWD/gadmm_b200/_native -o c_api_demo
 *
 * The packed layout is described in INTEGRATION.md: dp = 64*ceil(d/64), lower block triangle of
 * 64x64 row-major tiles, tile (I,J) at index I(I+1)/2+J, diagonal tiles stored full. */
#include <cuda_runtime_api.h>
#include <stdio.h>
#include <stdlib.h>

#include "gadmm_b200.h"

#define CHECK(call)                                                           \
  do {                                                                        \
    if ((call) != 0) {                                                        \
      fprintf(stderr, "%s failed: %s\n", #call, gadmm_last_error());          \
      return 1;                                                               \
    }                                                                         \
  } while (0)

int main(void) {
  const int N = 4, d = 100, dp = 128;                 /* four 100 x 100 matrices, padded to 128 */
  const size_t pe = gadmm_packed_elems(dp);
  double *dense = NULL, *packed = NULL, *x = NULL, *y = NULL;
  const void** ptrs = NULL;
  if (cudaMalloc((void**)&dense, sizeof(double) * N * dp * dp) != cudaSuccess ||
      cudaMalloc((void**)&packed, sizeof(double) * N * pe) != cudaSuccess ||
      cudaMalloc((void**)&x, sizeof(double) * N * dp) != cudaSuccess ||
      cudaMalloc((void**)&y, sizeof(double) * N * dp) != cudaSuccess ||
      cudaMalloc((void**)&ptrs, sizeof(void*) * N) != cudaSuccess) {
    fprintf(stderr, "cudaMalloc failed\n");
    return 1;
  }
  /* S_n = (n+1) * I on the host, x = ones: y must be (n+1) on the first d entries, 0 on the pad */
  double* h = (double*)calloc((size_t)N * dp * dp, sizeof(double));
  double* hx = (double*)calloc((size_t)N * dp, sizeof(double));
  const void* hp[4];
  for (int n = 0; n < N; ++n) {
    for (int i = 0; i < d; ++i) {
      h[((size_t)n * dp + i) * dp + i] = n + 1.0;
      hx[(size_t)n * dp + i] = 1.0;
    }
    hp[n] = packed + (size_t)n * pe;
  }
  cudaMemcpy(dense, h, sizeof(double) * N * dp * dp, cudaMemcpyHostToDevice);
  cudaMemcpy(x, hx, sizeof(double) * N * dp, cudaMemcpyHostToDevice);
  cudaMemcpy(ptrs, hp, sizeof(void*) * N, cudaMemcpyHostToDevice);

  CHECK(gadmm_pack_sym(dense, packed, N, d, dp, GADMM_STORE_F64, NULL));
  CHECK(gadmm_symv_batched_noalloc(ptrs, x, y, N, dp, GADMM_STORE_F64, NULL));
  cudaMemcpy(hx, y, sizeof(double) * N * dp, cudaMemcpyDeviceToHost);
  for (int n = 0; n < N; ++n)
    printf("matrix %d: y[0] = %g, y[%d] = %g, y[%d] (pad) = %g\n", n, hx[(size_t)n * dp], d - 1,
           hx[(size_t)n * dp + d - 1], dp - 1, hx[(size_t)n * dp + dp - 1]);
  free(h);
  free(hx);
  cudaFree(dense); cudaFree(packed); cudaFree(x); cudaFree(y); cudaFree((void*)ptrs);
  return 0;
}
