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

// ---------------------------------------------------------------------------------------------
// hvd_dense_forward_noalloc: the same Dense forward for callers that are not able to allocate
// device memory.
//
// Who needs it: code that runs where cudaMalloc is off limits - a plugin living inside another
// runtime's memory pool that forbids side allocations, a real-time / serving loop that must not
// take the allocator's device-wide lock (cudaMalloc synchronises the device), or a host that was
// handed a fixed memory budget at start-up.  Such a caller can still own x, W, bias and y, but it
// has nowhere to get the M*N split-K accumulator that hvd_dense_forward asks for.
//
// Here the accumulator is a fixed array inside this library's device image.  It costs
// HVD_DENSE_NOALLOC_MAX_ELEMS * 4 bytes (1 MiB) on EVERY GPU that loads libhvd_cuda.so, from load
// until the process exits, whether or not this function is ever called.
//
// Rules for callers:
//   * M*N must not exceed HVD_DENSE_NOALLOC_MAX_ELEMS; larger problems are refused with
//     HVD_ERR_NOALLOC_TOO_LARGE and a message from hvd_dense_forward_noalloc_error().
//   * The first call on each device looks up the array's device address (one-time set-up, with that
//     device current).  Make that first call before you begin any stream capture on the device.
//   * The array used is the one of the device that is CURRENT when you call; `stream` must belong to
//     that device (checked: a stream of another device is refused with HVD_ERR_NOALLOC_SETUP).
//   * All calls on a device share the one array and it keeps whatever the previous call left in it;
//     this function therefore clears the part it is about to use, on `stream`, before the GEMM.
//     Two calls whose GPU work can overlap (different streams, or different host threads without
//     ordering between their streams) race on the array and produce wrong results: serialise them.
//   * Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
// ---------------------------------------------------------------------------------------------
#include <mutex>
#include <string>

#define HVD_DENSE_NOALLOC_MAX_ELEMS (256 * 1024)
#define HVD_ERR_NOALLOC_TOO_LARGE (-100)
#define HVD_ERR_NOALLOC_SETUP (-101)

static __device__ __align__(256) float g_dense_noalloc_scratch[HVD_DENSE_NOALLOC_MAX_ELEMS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};     // device address of the array, per device ordinal
thread_local std::string g_noalloc_err;

// In host code `g_dense_noalloc_scratch` is only a stand-in symbol, not a device address: the real
// address has to be asked for, per device, with that device current.
int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    g_noalloc_err = std::string("cannot identify the current device: ") + cudaGetErrorString(e);
    return HVD_ERR_NOALLOC_SETUP;
  }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_dense_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      g_noalloc_err = std::string("cudaGetSymbolAddress(scratch) failed on device ") + std::to_string(dev) + ": " +
                      cudaGetErrorString(e);
      return HVD_ERR_NOALLOC_SETUP;
    }
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" const char* hvd_dense_forward_noalloc_error() { return g_noalloc_err.c_str(); }
extern "C" long long hvd_dense_forward_noalloc_max_elems() { return HVD_DENSE_NOALLOC_MAX_ELEMS; }

extern "C" int hvd_dense_forward_noalloc(const float* x, const float* W, const float* bias, float* y, int M, int N,
                                         int K, int relu, int k_split, void* stream) {
  g_noalloc_err.clear();
  if (M < 1 || N < 1 || K < 1 || k_split < 1) {
    g_noalloc_err = "M, N, K and k_split must be positive";
    return -1;
  }
  const long long need = (long long)M * N;
  if (need > HVD_DENSE_NOALLOC_MAX_ELEMS) {
    g_noalloc_err = "hvd_dense_forward_noalloc: M*N = " + std::to_string(need) + " output elements exceed the built-in "
                    "scratch of " + std::to_string((long long)HVD_DENSE_NOALLOC_MAX_ELEMS) +
                    " floats; use hvd_dense_forward with your own scratch buffer";
    return HVD_ERR_NOALLOC_TOO_LARGE;
  }
  if ((N & 3) || (K & 3)) {                // checked here so that nothing is enqueued for a call that is refused
    g_noalloc_err = "hvd_dense_forward_noalloc: N and K must be multiples of 4 (16-byte row strides for TMA)";
    return -2;
  }
  cudaStream_t st = (cudaStream_t)stream;
  // The array that is used is the CURRENT device's.  A stream that belongs to another device would make the
  // kernels run there with this device's address: refuse instead of touching the wrong GPU's memory.
  {
    int cur = -1, sdev = -1;
    if (cudaGetDevice(&cur) == cudaSuccess && cudaStreamGetDevice(st, &sdev) == cudaSuccess && sdev != cur) {
      g_noalloc_err = "hvd_dense_forward_noalloc: the stream belongs to device " + std::to_string(sdev) +
                      " but device " + std::to_string(cur) + " is current; make the stream's device current first";
      return HVD_ERR_NOALLOC_SETUP;
    }
    cudaGetLastError();                    // a failed query is not an error of this call
  }
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  // shared array, previous contents unknown: the accumulator region must start at zero
  cudaError_t e = cudaMemsetAsync(scratch, 0, (size_t)need * sizeof(float), st);
  if (e != cudaSuccess) {
    g_noalloc_err = std::string("clearing the scratch failed: ") + cudaGetErrorString(e);
    return (int)e;
  }
  rc = hvd_dense_forward_impl(x, W, bias, y, M, N, K, relu, k_split, scratch, st);
  if (rc != 0)
    g_noalloc_err = "hvd_dense_forward_noalloc: the GEMM or its finalize pass failed with code " + std::to_string(rc) +
                    (rc > 0 ? std::string(" (") + cudaGetErrorString((cudaError_t)rc) + ")" : std::string(" (tensor-map or argument error)"));
  return rc;
}
