// conv2 forward: Conv2D(64, 3x3, relu) + MaxPooling2D(2x2) + Dropout(0.25) + Flatten in ONE kernel
// (reference layers: mnist_keras.py:75-78).  Implicit GEMM on tcgen05 / TMEM fed by TMA:
//
//   M = output pixels (B*24*24 = 73,728), N = 64 output channels, K = 9 taps x 32 input channels.
//   An M tile is two 8x8 spatial patches (128 pixels), so every 2x2 pool window lies inside one
//   tile and inside one warp's 32 TMEM lanes: pooling is two __shfl_xor per value in the epilogue.
//   A (im2col) is never materialised: for tap (kh,kw) the K-slice of a patch is the 4-D TMA box
//   {32 ch, 8 w, 8 h, 1 n} of a1 at (0, ow0+kw, oh0+kh, n) - 64 rows of 128 B, 128B-swizzled.
//   B = the whole HWIO kernel (72 KB) stays resident in smem for the life of the persistent CTA;
//   per tap it is [32 cin][64 cout] with cout contiguous = an MN-major operand (2 atoms of 32).
//
// Persistent, warp-specialised: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator,
// warps 4-7 epilogue.  Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of
// tile i+1.  The epilogue writes only the pooled, dropped-out activations (4.7 MB instead of the
// 18.9 MB conv output) plus one mask byte per element (arg-max + keep) for backward.
#include "mnist_common.cuh"
#include "tc_common.cuh"

namespace {
using namespace mnist;

constexpr int STAGES = 4;
constexpr int A_STAGE = 128 * 128;          // 2 patches x 64 rows x 128 B
constexpr int W_BYTES = 9 * 8192;           // 9 taps x [32 cin][64 cout] fp32
constexpr int SMEM_BYTES = W_BYTES + STAGES * A_STAGE + 1024 + 256;
constexpr int TMEM_COLS = 128;              // 2 accumulator stages x 64 columns

struct Params {
  int B, n_tiles, training;
  float keep;                               // 1 - dropout rate
  const float* bias;
  float* p2;                                // [B, 12, 12, 64]
  uint8_t* mask;                            // [B, 12, 12, 64]
  const StepState* st;
};

__global__ void __launch_bounds__(256, 1)
conv2_fprop_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem;
  uint8_t* sA = smem + W_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sA + STAGES * A_STAGE);
  uint64_t* full = bars;               // [STAGES]
  uint64_t* empty = bars + STAGES;     // [STAGES]
  uint64_t* wfull = bars + 2 * STAGES;
  uint64_t* accf = wfull + 1;          // [2] accumulator full
  uint64_t* acce = accf + 2;           // [2] accumulator drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acce + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) { tc::tma_prefetch_desc(&tmA); tc::tma_prefetch_desc(&tmW); }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    tc::mbar_init(wfull, 1);
    for (int s = 0; s < 2; s++) { tc::mbar_init(&accf[s], 1); tc::mbar_init(&acce[s], 128); }
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc<TMEM_COLS>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int n_patches = p.B * 9;

  if (warp == 0) {
    if (tc::elect_one()) {
      tc::mbar_expect_tx(wfull, W_BYTES);
      for (int t = 0; t < 9; t++)
        for (int j = 0; j < 2; j++) tc::tma_load_2d(sW + t * 8192 + j * 4096, &tmW, wfull, 32 * j, t * 32);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        for (int tap = 0; tap < 9; tap++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          tc::mbar_expect_tx(&full[s], A_STAGE);
          const int kh = tap / 3, kw = tap % 3;
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int patch = 2 * tile + h;        // patch >= n_patches reads n >= B: TMA zero-fills
            const int b = patch / 9, py = (patch % 9) / 3, px = patch % 3;
            tc::tma_load_4d(sA + s * A_STAGE + h * 8192, &tmA, &full[s], 0, px * 8 + kw, py * 8 + kh, b);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(128, 64, 0, 1);
      tc::mbar_wait(wfull, 0);
      int it = 0, ti = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ti++) {
        const int as = ti & 1;
        tc::mbar_wait(&acce[as], ((ti >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        for (int tap = 0; tap < 9; tap++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&full[s], (it / STAGES) & 1);
          tc::tc_fence_after();
          const uint32_t a = tc::smem_u32(sA + s * A_STAGE);
          const uint32_t w = tc::smem_u32(sW + tap * 8192);
#pragma unroll
          for (int k = 0; k < 4; k++)
            tc::mma_tf32(tmem + as * 64, tc::desc_kmajor(a + k * 32), tc::desc_mnmajor(w + k * 1024, 4096), idesc,
                         (tap | k) != 0);
          tc::mma_commit(&empty[s]);
        }
        tc::mma_commit(&accf[as]);
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int r = q * 32 + lane;                 // accumulator row = pixel inside the tile
    const int half = r >> 6, hl = (r & 63) >> 3, wl = r & 7;
    const bool owner = (lane & 9) == 0;          // even h, even w: writes the pooled element
    const unsigned step = p.st->step, seed = p.st->seed;
    const float inv_keep = 1.0f / p.keep;
    int ti = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ti++) {
      const int as = ti & 1;
      tc::mbar_wait(&accf[as], (ti >> 1) & 1);
      tc::tc_fence_after();
      float v[2][32];
      tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + as * 64, v[0]);
      tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + as * 64 + 32, v[1]);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acce[as]);                // TMEM stage is free: next tile's MMAs may start
      const int patch = 2 * tile + half;
      const int b = patch / 9, py = (patch % 9) / 3, px = patch % 3;
      const int ph = py * 4 + (hl >> 1), pw = px * 4 + (wl >> 1);
      const long long obase = (((long long)b * HP + ph) * HP + pw) * C2;
#pragma unroll
      for (int c = 0; c < 2; c++) {
        unsigned flags = 0;
        float m[32];
#pragma unroll
        for (int j = 0; j < 32; j++) {
          const float x = fmaxf(v[c][j] + __ldg(p.bias + c * 32 + j), 0.f);
          float mx = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
          m[j] = mx;
          flags |= (x == mx ? 1u : 0u) << j;
        }
        const unsigned f1 = __shfl_xor_sync(0xffffffffu, flags, 1);
        const unsigned f8 = __shfl_xor_sync(0xffffffffu, flags, 8);
        if (owner && patch < n_patches) {
          uint32_t mb[8];
#pragma unroll
          for (int j4 = 0; j4 < 8; j4++) {
            uint4 rnd = make_uint4(0, 0, 0, 0);
            if (p.training) rnd = philox4x32_10(make_uint4((unsigned)((obase + c * 32) >> 2) + j4, step, 0x2u, 0u), make_uint2(seed, 0x5EEDu));
            const unsigned rr[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
            float o[4];
            uint32_t packed = 0;
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const int j = j4 * 4 + e;
              const unsigned pos = ((flags >> j) & 1) ? 0u : ((f1 >> j) & 1) ? 1u : ((f8 >> j) & 1) ? 2u : 3u;
              const bool keep = !p.training || u01(rr[e]) < p.keep;
              const float val = keep ? (p.training ? m[j] * inv_keep : m[j]) : 0.f;
              o[e] = val;
              packed |= (pos | ((keep && m[j] > 0.f) ? MASK_LIVE : 0u)) << (8 * e);
            }
            mb[j4] = packed;
            *reinterpret_cast<float4*>(p.p2 + obase + c * 32 + j4 * 4) = make_float4(o[0], o[1], o[2], o[3]);
          }
          uint4* mp = reinterpret_cast<uint4*>(p.mask + obase + c * 32);
          mp[0] = make_uint4(mb[0], mb[1], mb[2], mb[3]);
          mp[1] = make_uint4(mb[4], mb[5], mb[6], mb[7]);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc<TMEM_COLS>(tmem);
}

}  // namespace

// a1: [B,26,26,32] fp32 (conv1 output, post-relu); w2: HWIO [3,3,32,64]; bias2[64];
// p2: [B,12,12,64]; mask: [B,12,12,64] bytes.  Returns 0 or a CUDA/driver error code.
extern "C" int hvd_conv2_fprop(const float* a1, const float* w2, const float* bias2, float* p2, uint8_t* mask,
                               const void* step_state, int B, int training, float keep, int ctas, void* stream) {
  CUtensorMap tmA, tmW;
  {
    uint64_t dims[4] = {32, 26, 26, (uint64_t)B};
    uint64_t str[3] = {128, 26 * 128, 26 * 26 * 128};
    uint32_t box[4] = {32, 8, 8, 1};
    if (tc::make_tmap(&tmA, a1, 4, dims, str, box)) return -2;
    uint64_t wd[2] = {64, 288};
    uint64_t ws[1] = {256};
    uint32_t wb[2] = {32, 32};
    if (tc::make_tmap(&tmW, w2, 2, wd, ws, wb, true)) return -3;
  }
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(conv2_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) return -10;
    attr = true;
  }
  Params p;
  p.B = B; p.n_tiles = (B * 9 + 1) / 2; p.training = training; p.keep = keep; p.bias = bias2; p.p2 = p2; p.mask = mask;
  p.st = reinterpret_cast<const mnist::StepState*>(step_state);
  if (ctas <= 0) ctas = 148;
  if (ctas > p.n_tiles) ctas = p.n_tiles;
  conv2_fprop_kernel<<<ctas, 256, SMEM_BYTES, (cudaStream_t)stream>>>(tmA, tmW, p);
  return (int)cudaGetLastError();
}
