// Dense layer forward  y = act(x * W + bias)  as a split-K tcgen05 GEMM + finalize pass
// (keras.layers.Dense; reference mnist_keras.py:79,81).
//
// With batch 128 the output is a single 128 x N tile, so the K dimension (9216 for the reference's
// Dense(128)) is what gets split across CTAs; each slice adds its partial tile into a scratch
// accumulator with red.add and a small second kernel applies bias + activation and re-arms
// (zeroes) the accumulator for the next call.
#include <cuda_runtime.h>

#include "hvd_b200.h"

namespace {

__global__ void __launch_bounds__(256) dense_finalize_kernel(float* __restrict__ scratch, const float* __restrict__ bias,
                                                             float* __restrict__ y, int M, int N, int relu) {
  const long long total = (long long)M * N;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    float v = scratch[i] + (bias ? bias[i % N] : 0.f);
    scratch[i] = 0.f;
    y[i] = relu ? fmaxf(v, 0.f) : v;
  }
}

}  // namespace

int hvd_dense_forward_impl(const float* x, const float* W, const float* bias, float* y, int M, int N, int K, int relu,
                           int k_split, float* scratch, cudaStream_t st) {
  if (M < 1 || N < 1 || K < 1 || k_split < 1 || !scratch) return -1;
  if ((N & 3) || (K & 3)) return -2;     // TMA needs 16-byte row strides: K and N multiples of 4
  const int bn = (N % 128 == 0 || N > 64) ? 128 : 64;
  int rc = hvd_tc_gemm(x, W, scratch, M, N, K, /*a_mn=*/0, /*b_mn=*/1, bn, k_split, /*atomic add*/ 1, nullptr, N, st,
                       0, 0, 0, 0, nullptr, 1.0f);
  if (rc != 0) return rc;
  const long long total = (long long)M * N;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 592) blocks = 592;
  dense_finalize_kernel<<<blocks, 256, 0, st>>>(scratch, bias, y, M, N, relu);
  return (int)cudaGetLastError();
}

extern "C" int hvd_dense_forward(const float* x, const float* W, const float* bias, float* y, int M, int N, int K,
                                 int relu, int k_split, float* scratch, void* stream) {
  return hvd_dense_forward_impl(x, W, bias, y, M, N, K, relu, k_split, scratch, (cudaStream_t)stream);
}
