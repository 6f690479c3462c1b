// Shared definitions for the fused MNIST-CNN engine kernels (reference model:
// mnist_keras.py:71-81 / tensorflow2_keras_mnist.py:43-52).
//
// Shapes (NHWC, batch B):  x[B,28,28,1] -> conv1 -> a1[B,26,26,32] -> conv2 -> [B,24,24,64]
//   -> maxpool 2x2 + dropout 0.25 -> p2[B,12,12,64] == flatten [B,9216] (h,w,c order, as Keras)
//   -> dense 128 relu dropout 0.5 -> dense 10 softmax.
// Weights keep Keras' on-disk layouts: conv kernels HWIO, dense kernels (in,out).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mnist {

constexpr int IMG = 28, C1 = 32, H1 = 26, C2 = 64, H2 = 24, HP = 12, FLAT = HP * HP * C2 /*9216*/, HID = 128, NCLS = 10;

// device-side step state shared by the kernels of one training step
struct StepState {
  unsigned int step;        // incremented once per training step (by the head kernel)
  float loss_sum;           // running sum of per-batch mean losses since last reset
  float correct_sum;        // running count of correct predictions
  float last_loss;          // mean loss of the latest batch
  float last_correct;       // correct count of the latest batch
  unsigned int seed;
  unsigned int pad0, pad1;
};

// Philox4x32-10 (Salmon et al.), the counter-based RNG TF/Keras dropout is built on.  Masks
// are a pure function of (seed, step, element), so backward could regenerate them; we store
// the keep bit anyway because it travels with the pool arg-max in one byte.
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; i++) {
    const unsigned int hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const unsigned int hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
__device__ __forceinline__ float u01(unsigned int r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }

// mask byte written by the conv2 epilogue for every pooled element
//   bits 0-1 : arg-max position inside the 2x2 window (dy*2+dx)
//   bit  2   : gradient flows (kept by dropout AND relu output > 0)
constexpr uint8_t MASK_LIVE = 4;

}  // namespace mnist
