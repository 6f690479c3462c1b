// conv2 backward on tcgen05 / TMEM / TMA (reference layer: Conv2D(64,(3,3)) at mnist_keras.py:75).
//
// dgrad  da1[b,y,x,ci] = relu'(a1) * sum_{kh,kw,co} dY2[b,y-kh,x-kw,co] * W2[kh,kw,ci,co]
//        implicit GEMM: M = 26x26 output pixels in 13x9 boxes (117 of 128 rows used), N = 32 ci,
//        K = 9 taps x 64 co in 18 blocks of 32.  A = 4-D TMA boxes of dY2 at NEGATIVE offsets - the
//        "full" convolution's zero padding is TMA out-of-bounds fill, no padded copy exists.
//        B = W2[tap] as [ci][co] (co contiguous = K-major), whole kernel resident in smem.
//        Epilogue applies conv1's ReLU mask and stores da1.
//
// wgrad  dW2[kh,kw,ci,co] = sum_{b,oy,ox} a1[b,oy+kh,ox+kw,ci] * dY2[b,oy,ox,co]
//        the reduction runs over pixels, so BOTH operands are MN-major: A[(tap,ci)][pixel] and
//        B[co][pixel] are the activation tensors themselves ([pixel][channel] rows).  M = 4 taps x
//        32 ci per accumulator, three accumulators in TMEM (taps 0-3, 4-7, 8 + a tile of ones whose
//        "tap" yields the bias gradient for free).  Each CTA reduces a slice of the pixels
//        (split-K over 8x8 patches) and adds its 288x64 result into the arena with red.add.
#include "mnist_common.cuh"
#include "tc_common.cuh"

namespace {
using namespace mnist;

// ============================================================================ dgrad
namespace dg {
constexpr int STAGES = 6;
constexpr int A_STAGE = 128 * 128;
constexpr int A_TX = 117 * 128;             // bytes one 13x9 box delivers
constexpr int W_BYTES = 18 * 4096;
constexpr int SMEM_BYTES = W_BYTES + STAGES * A_STAGE + 1024 + 256;
constexpr int TMEM_COLS = 64;               // 2 stages x 32 columns
constexpr int BW = 13, BH = 9;              // box: 13 wide, 9 high -> 2 x 3 boxes per 26x26 image

struct Params {
  int B, n_tiles;
  const float* a1;
  float* da1;
};

__global__ void __launch_bounds__(256, 1)
conv2_dgrad_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmW, Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem;
  uint8_t* sA = smem + W_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sA + STAGES * A_STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* wfull = bars + 2 * STAGES;
  uint64_t* accf = wfull + 1;
  uint64_t* acce = accf + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acce + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) { tc::tma_prefetch_desc(&tmG); tc::tma_prefetch_desc(&tmW); }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    tc::mbar_init(wfull, 1);
    for (int s = 0; s < 2; s++) { tc::mbar_init(&accf[s], 1); tc::mbar_init(&acce[s], 128); }
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc<TMEM_COLS>(tmem_slot);
  // rows 117..127 of every A stage are never written by TMA: zero them once so no NaN bit
  // patterns ever enter the tensor core (their accumulator rows are discarded anyway)
  for (int i = threadIdx.x; i < STAGES * 11 * 32; i += blockDim.x) {
    const int s = i / (11 * 32), r = i % (11 * 32);
    reinterpret_cast<float*>(sA + s * A_STAGE + 117 * 128)[r] = 0.f;
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (tc::elect_one()) {
      tc::mbar_expect_tx(wfull, W_BYTES);
      for (int t = 0; t < 9; t++)
        for (int h = 0; h < 2; h++) tc::tma_load_2d(sW + (t * 2 + h) * 4096, &tmW, wfull, 32 * h, t * 32);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const int b = tile / 6, ty = (tile % 6) / 2, tx = tile % 2;
        for (int kb = 0; kb < 18; kb++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          tc::mbar_expect_tx(&full[s], A_TX);
          const int tap = kb >> 1, kh = tap / 3, kw = tap % 3;
          tc::tma_load_4d(sA + s * A_STAGE, &tmG, &full[s], 32 * (kb & 1), tx * BW - kw, ty * BH - kh, b);
        }
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(128, 32, 0, 0);
      tc::mbar_wait(wfull, 0);
      int it = 0, ti = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ti++) {
        const int as = ti & 1;
        tc::mbar_wait(&acce[as], ((ti >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        for (int kb = 0; kb < 18; kb++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&full[s], (it / STAGES) & 1);
          tc::tc_fence_after();
          const uint32_t a = tc::smem_u32(sA + s * A_STAGE);
          const uint32_t w = tc::smem_u32(sW + kb * 4096);
#pragma unroll
          for (int k = 0; k < 4; k++)
            tc::mma_tf32(tmem + as * 32, tc::desc_kmajor(a + k * 32), tc::desc_kmajor(w + k * 32), idesc, (kb | k) != 0);
          tc::mma_commit(&empty[s]);
        }
        tc::mma_commit(&accf[as]);
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int r = q * 32 + lane;
    int ti = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ti++) {
      const int as = ti & 1;
      tc::mbar_wait(&accf[as], (ti >> 1) & 1);
      tc::tc_fence_after();
      float v[32];
      tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + as * 32, v);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acce[as]);
      const int b = tile / 6, ty = (tile % 6) / 2, tx = tile % 2;
      const int y = ty * BH + r / BW, x = tx * BW + r % BW;
      if (r < BW * BH && y < H1) {
        const long long off = (((long long)b * H1 + y) * H1 + x) * C1;
        const float4* act = reinterpret_cast<const float4*>(p.a1 + off);
        float4* dst = reinterpret_cast<float4*>(p.da1 + off);
#pragma unroll
        for (int j = 0; j < 8; j++) {
          const float4 a = __ldg(act + j);
          dst[j] = make_float4(a.x > 0.f ? v[4 * j] : 0.f, a.y > 0.f ? v[4 * j + 1] : 0.f,
                               a.z > 0.f ? v[4 * j + 2] : 0.f, a.w > 0.f ? v[4 * j + 3] : 0.f);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc<TMEM_COLS>(tmem);
}
}  // namespace dg

// ============================================================================ wgrad
namespace wg {
constexpr int SLOT = 8192;                  // one 8x8-pixel box of 32 channels
constexpr int STAGE = 12 * SLOT;            // 9 taps | ones | dY2 lo | dY2 hi
constexpr int STAGE_TX = 11 * SLOT;         // the ones slot is not loaded
constexpr int STAGES = 2;
constexpr int SMEM_BYTES = STAGES * STAGE + 1024 + 256;
constexpr int TMEM_COLS = 256;              // 3 accumulators x 64 columns

struct Params {
  int B, n_patches, patches_per_cta;
  float* g_w;                               // [288][64]
  float* g_b;                               // [64]
};

__global__ void __launch_bounds__(256, 1)
conv2_wgrad_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmG, Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* accf = bars + 2 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accf + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p_begin = blockIdx.x * p.patches_per_cta;
  const int p_end = min(p.n_patches, p_begin + p.patches_per_cta);
  const int np = p_end - p_begin;
  if (warp == 0 && lane == 0) { tc::tma_prefetch_desc(&tmA); tc::tma_prefetch_desc(&tmG); }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    tc::mbar_init(accf, 1);
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc<TMEM_COLS>(tmem_slot);
  for (int i = threadIdx.x; i < STAGES * (SLOT / 4); i += blockDim.x) {       // the tile of ones (bias gradient)
    const int s = i / (SLOT / 4), r = i % (SLOT / 4);
    reinterpret_cast<float*>(smem + s * STAGE + 9 * SLOT)[r] = 1.0f;
  }
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (tc::elect_one()) {
      for (int i = 0; i < np; i++) {
        const int s = i % STAGES;
        tc::mbar_wait(&empty[s], ((i / STAGES) & 1) ^ 1);
        tc::mbar_expect_tx(&full[s], STAGE_TX);
        const int patch = p_begin + i;
        const int b = patch / 9, py = (patch % 9) / 3, px = patch % 3;
        uint8_t* st = smem + s * STAGE;
        for (int t = 0; t < 9; t++) tc::tma_load_4d(st + t * SLOT, &tmA, &full[s], 0, px * 8 + t % 3, py * 8 + t / 3, b);
        tc::tma_load_4d(st + 10 * SLOT, &tmG, &full[s], 0, px * 8, py * 8, b);
        tc::tma_load_4d(st + 11 * SLOT, &tmG, &full[s], 32, px * 8, py * 8, b);
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(128, 64, 1, 1);
      for (int i = 0; i < np; i++) {
        const int s = i % STAGES;
        tc::mbar_wait(&full[s], (i / STAGES) & 1);
        tc::tc_fence_after();
        const uint32_t st = tc::smem_u32(smem + s * STAGE);
#pragma unroll
        for (int g = 0; g < 3; g++)
#pragma unroll
          for (int ks = 0; ks < 8; ks++)     // 64 pixels = 8 MMAs of K=8
            tc::mma_tf32(tmem + g * 64, tc::desc_mnmajor(st + g * 4 * SLOT + ks * 1024, SLOT),
                         tc::desc_mnmajor(st + 10 * SLOT + ks * 1024, SLOT), idesc, (i | ks) != 0);
        tc::mma_commit(&empty[s]);
      }
      tc::mma_commit(accf);
    }
  } else if (warp >= 4 && np > 0) {
    const int q = warp & 3;
    const int r = q * 32 + lane;              // accumulator row: tap-in-group = r/32, ci = r%32
    tc::mbar_wait(accf, 0);
    tc::tc_fence_after();
#pragma unroll 1
    for (int g = 0; g < 3; g++) {
      const int tap = g * 4 + (r >> 5);
      if (g == 2 && q >= 2) break;            // rows 64..127 of the last accumulator are padding
#pragma unroll 1
      for (int c = 0; c < 64; c += 32) {
        float v[32];
        tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + g * 64 + c, v);
        tc::tmem_ld_wait();
        if (tap < 9) {
          float* d = p.g_w + ((long long)tap * 32 + (r & 31)) * 64 + c;
#pragma unroll
          for (int j = 0; j < 32; j += 4) tc::red_add_f4(d + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else if (tap == 9 && (r & 31) == 0) {   // the ones "tap": every row holds sum_pixels dY2[co]
#pragma unroll
          for (int j = 0; j < 32; j++) atomicAdd(p.g_b + c + j, v[j]);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc<TMEM_COLS>(tmem);
}
}  // namespace wg

}  // namespace

extern "C" int hvd_conv2_dgrad(const float* dy2, const float* w2, const float* a1, float* da1, int B, int ctas, void* stream) {
  CUtensorMap tmG, tmW;
  {
    uint64_t dims[4] = {64, 24, 24, (uint64_t)B};
    uint64_t str[3] = {256, 24 * 256, 24 * 24 * 256};
    uint32_t box[4] = {32, dg::BW, dg::BH, 1};
    if (tc::make_tmap(&tmG, dy2, 4, dims, str, box)) return -2;
    uint64_t wd[2] = {64, 288};
    uint64_t ws[1] = {256};
    uint32_t wb[2] = {32, 32};
    if (tc::make_tmap(&tmW, w2, 2, wd, ws, wb)) return -3;
  }
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(dg::conv2_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dg::SMEM_BYTES) != cudaSuccess) return -10;
    attr = true;
  }
  dg::Params p{B, B * 6, a1, da1};
  if (ctas <= 0) ctas = 148;
  if (ctas > p.n_tiles) ctas = p.n_tiles;
  dg::conv2_dgrad_kernel<<<ctas, 256, dg::SMEM_BYTES, (cudaStream_t)stream>>>(tmG, tmW, p);
  return (int)cudaGetLastError();
}

// g_w / g_b are ACCUMULATED into (red.add): the caller zeroes them before the step.
extern "C" int hvd_conv2_wgrad(const float* a1, const float* dy2, float* g_w, float* g_b, int B, int ctas, void* stream) {
  CUtensorMap tmA, tmG;
  {
    uint64_t dims[4] = {32, 26, 26, (uint64_t)B};
    uint64_t str[3] = {128, 26 * 128, 26 * 26 * 128};
    uint32_t box[4] = {32, 8, 8, 1};
    if (tc::make_tmap(&tmA, a1, 4, dims, str, box, true)) return -2;
    uint64_t gd[4] = {64, 24, 24, (uint64_t)B};
    uint64_t gs[3] = {256, 24 * 256, 24 * 24 * 256};
    if (tc::make_tmap(&tmG, dy2, 4, gd, gs, box, true)) return -3;
  }
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(wg::conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wg::SMEM_BYTES) != cudaSuccess) return -10;
    attr = true;
  }
  wg::Params p;
  p.B = B; p.n_patches = B * 9;
  if (ctas <= 0) ctas = 72;
  p.patches_per_cta = (p.n_patches + ctas - 1) / ctas;
  ctas = (p.n_patches + p.patches_per_cta - 1) / p.patches_per_cta;
  p.g_w = g_w; p.g_b = g_b;
  wg::conv2_wgrad_kernel<<<ctas, 256, wg::SMEM_BYTES, (cudaStream_t)stream>>>(tmA, tmG, p);
  return (int)cudaGetLastError();
}
