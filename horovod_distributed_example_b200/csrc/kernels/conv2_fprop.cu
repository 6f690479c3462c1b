// conv2 forward: Conv2D(64, 3x3, relu) + MaxPooling2D(2x2) + Dropout(0.25) + Flatten in ONE kernel
// (reference layers: mnist_keras.py:75-78).  Implicit GEMM on tcgen05 / TMEM fed by TMA:
//
//   M = output pixels (B*24*24 = 73,728), N = 64 output channels, K = 9 taps x 32 input channels.
//   An M tile is the same 8x8 spatial patch of two consecutive images (128 pixels), so every 2x2
//   pool window lies inside one tile and inside one warp's 32 TMEM lanes: pooling is two __shfl_xor
//   per value in the epilogue.
//   A (im2col) is never materialised, and the 9 taps do not re-read their inputs: per kernel row kh
//   ONE 4-D TMA box {32 ch, 10 w, 8 h, 2 n} brings a 10-wide halo (160 rows of 128 B, 128B-swizzled)
//   and the three kw taps are the same smem tile read through descriptors whose start is shifted by
//   kw rows with a group stride of 10 rows (SBO = 1280 B).  The tensor core swizzles on absolute
//   address bits, so shifted starts and non-power-of-two strides are legal (profiles/
//   desc_probe_shifted_descriptors.json).  2.45x less L2->smem traffic than one box per tap.
//   B = the whole HWIO kernel (72 KB) stays resident in smem for the life of the persistent CTA;
//   per tap it is [32 cin][64 cout] with cout contiguous = an MN-major operand (2 atoms of 32).
//
// Persistent, warp-specialised: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator,
// 16 epilogue warps in two sets.  Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of
// tile i+1.  The epilogue writes only the pooled, dropped-out activations (4.7 MB instead of the
// 18.9 MB conv output) plus one mask byte per element (arg-max + keep) for backward.
#include "comm/pipeline.cuh"
#include "mnist_common.cuh"
#include "tc_common.cuh"

namespace {
using namespace mnist;

constexpr int STAGES = 4;
constexpr int A_STAGE = 20 * 1024;          // one kh: box {32 c, 10 w, 8 h, 2 n} = 160 rows x 128 B
constexpr int W_BYTES = 9 * 8192;           // 9 taps x [32 cin][64 cout] fp32
constexpr int SMEM_BYTES = W_BYTES + STAGES * A_STAGE + 1024 + 256;
constexpr int TMEM_COLS = 128;              // 2 accumulator stages x 64 columns
constexpr int EPI_SETS = 2;                 // warp sets: set s drains accumulator stage s (tiles alternate)
constexpr int EPI_WARPS = 8 * EPI_SETS;     // per set: 2 warps per TMEM lane quadrant (one per 32-column chunk)
constexpr int THREADS = 128 + EPI_WARPS * 32;
constexpr uint32_t SBO_HALO = 10 * 128;     // next output row = 10 halo rows further

struct Params {
  int B, n_tiles, training;
  float keep;                               // 1 - dropout rate
  const float* bias;
  float* p2;                                // [B, 12, 12, 64]
  uint8_t* mask;                            // [B, 12, 12, 64]
  const StepState* st;
  // Pipelined data-parallel step: the spare warp of CTA 0 waits here until every peer has delivered its rows of
  // dense_1/kernel (comm/pipeline.cuh).  This kernel does not read those weights - the wait rides along because
  // it costs several microseconds of system-scope polling and fencing that would otherwise sit in front of the
  // Dense forward; the kernel boundary orders the Dense forward's reads after it.
  HvdParamWait wait;
};

__global__ void __launch_bounds__(THREADS, 1)
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
    for (int s = 0; s < 2; s++) { tc::mbar_init(&accf[s], 1); tc::mbar_init(&acce[s], 8 * 32); }
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc<TMEM_COLS>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  tc::pdl_launch_dependents();
  tc::pdl_wait();

  if (warp == 0) {
    if (tc::elect_one()) {
      tc::mbar_expect_tx(wfull, W_BYTES);
      for (int t = 0; t < 9; t++)
        for (int j = 0; j < 2; j++) tc::tma_load_2d(sW + t * 8192 + j * 4096, &tmW, wfull, 32 * j, t * 32);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const int ip = tile / 9, py = (tile % 9) / 3, px = tile % 3;
        for (int kh = 0; kh < 3; kh++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          tc::mbar_expect_tx(&full[s], A_STAGE);
          // rows kh..kh+7 of the 10-wide halo of BOTH images of the pair (image >= B: TMA zero-fills)
          tc::tma_load_4d(sA + s * A_STAGE, &tmA, &full[s], 0, px * 8, py * 8 + kh, 2 * ip);
        }
      }
    }
  } else if (warp == 3) {
    if (blockIdx.x == 0) hvd_param_wait_warp(p.wait);
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(128, 64, 0, 1);
      const uint64_t dW0 = tc::desc_mnmajor(tc::smem_u32(sW), 4096);
      tc::mbar_wait(wfull, 0);
      int it = 0, ti = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ti++) {
        const int as = ti & 1;
        tc::mbar_wait(&acce[as], ((ti >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        for (int kh = 0; kh < 3; kh++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&full[s], (it / STAGES) & 1);
          tc::tc_fence_after();
          const uint64_t da0 = tc::make_smem_desc(tc::smem_u32(sA + s * A_STAGE), 16, SBO_HALO, 2);
          const uint64_t dw0 = tc::desc_advance(dW0, kh * 3 * 8192);
#pragma unroll
          for (int kw = 0; kw < 3; kw++) {
#pragma unroll
            for (int k = 0; k < 4; k++)     // the tap shift is a row offset into the shared halo tile
              tc::mma_tf32(tmem + as * 64, tc::desc_advance(da0, kw * 128 + k * 32),
                           tc::desc_advance(dw0, kw * 8192 + k * 1024), idesc, (kh | kw | k) != 0);
          }
          tc::mma_commit(&empty[s]);
        }
        tc::mma_commit(&accf[as]);
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;                      // TMEM lane quadrant
    const int c = ((warp - 4) >> 2) & 1;         // which 32-channel chunk this warp handles
    const int wset = (warp - 4) >> 3;            // warp set 0 drains even tiles (stage 0), set 1 odd tiles
    const int r = q * 32 + lane;                 // accumulator row = pixel inside the tile
    const int half = r >> 6, hl = (r & 63) >> 3, wl = r & 7;
    const int wp = (lane & 1) | ((lane >> 2) & 2);   // position inside the 2x2 window (dy*2+dx)
    const int base = lane & ~9;
    const unsigned step = p.st->step, seed = p.st->seed;
    const float inv_keep = 1.0f / p.keep;
    float bias[32];
#pragma unroll
    for (int j = 0; j < 32; j++) bias[j] = __ldg(p.bias + c * 32 + j);
    int ti = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ti++) {
      const int as = ti & 1;
      if (EPI_SETS == 2 && as != wset) continue;
      tc::mbar_wait(&accf[as], (ti >> 1) & 1);
      tc::tc_fence_after();
      float v[32];
      tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + as * 64 + c * 32, v);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acce[as]);                // TMEM stage is free: next tile's MMAs may start
      const int ip = tile / 9, py = (tile % 9) / 3, px = tile % 3;
      const int b = 2 * ip + half;
      const int ph = py * 4 + (hl >> 1), pw = px * 4 + (wl >> 1);
      const long long obase = (((long long)b * HP + ph) * HP + pw) * C2 + c * 32;
      unsigned flags = 0;
#pragma unroll
      for (int j = 0; j < 32; j++) {
        const float x = fmaxf(v[j] + bias[j], 0.f);
        float mx = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
        v[j] = mx;                               // every lane of the window now holds the window max
        flags |= (x == mx ? 1u : 0u) << j;
      }
      // arg-max bookkeeping in the window's canonical order (pos 0,1,2; else 3)
      const unsigned f0 = __shfl_sync(0xffffffffu, flags, base);
      const unsigned f1 = __shfl_sync(0xffffffffu, flags, base | 1);
      const unsigned f2 = __shfl_sync(0xffffffffu, flags, base | 8);
      if (b < p.B) {
        // the four lanes of a window split its 32 channels: lane `wp` finishes channels 8*wp..8*wp+7
        uint32_t mb[2];
        float o[8];
#pragma unroll
        for (int h4 = 0; h4 < 2; h4++) {
          uint4 rnd = make_uint4(0, 0, 0, 0);
          if (p.training)
            rnd = philox4x32_10(make_uint4((unsigned)((obase + 8 * wp) >> 2) + h4, step, 0x2u, 0u), make_uint2(seed, 0x5EEDu));
          const unsigned rr[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
          uint32_t packed = 0;
#pragma unroll
          for (int e = 0; e < 4; e++) {
            const int jj = h4 * 4 + e;           // 0..7 within this lane's slice
            float m = 0.f;
            unsigned b0 = 0, b1 = 0, b2 = 0;
#pragma unroll
            for (int w4 = 0; w4 < 4; w4++)       // select channel 8*wp+jj without dynamic register indexing
              if (w4 == wp) {
                m = v[w4 * 8 + jj];
                b0 = (f0 >> (w4 * 8 + jj)) & 1; b1 = (f1 >> (w4 * 8 + jj)) & 1; b2 = (f2 >> (w4 * 8 + jj)) & 1;
              }
            const unsigned pos = b0 ? 0u : b1 ? 1u : b2 ? 2u : 3u;
            const bool keep = !p.training || u01(rr[e]) < p.keep;
            o[jj] = keep ? (p.training ? m * inv_keep : m) : 0.f;
            packed |= (pos | ((keep && m > 0.f) ? MASK_LIVE : 0u)) << (8 * e);
          }
          mb[h4] = packed;
        }
        float4* dst = reinterpret_cast<float4*>(p.p2 + obase + 8 * wp);
        dst[0] = make_float4(o[0], o[1], o[2], o[3]);
        dst[1] = make_float4(o[4], o[5], o[6], o[7]);
        *reinterpret_cast<uint2*>(p.mask + obase + 8 * wp) = make_uint2(mb[0], mb[1]);
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
extern "C" int hvd_conv2_fprop_ex(const float* a1, const float* w2, const float* bias2, float* p2, uint8_t* mask,
                                  const void* step_state, int B, int training, float keep, int ctas, void* stream,
                                  const HvdParamWait* wait) {
  CUtensorMap tmA, tmW;
  {
    uint64_t dims[4] = {32, 26, 26, (uint64_t)B};
    uint64_t str[3] = {128, 26 * 128, 26 * 26 * 128};
    uint32_t box[4] = {32, 10, 8, 2};     // 10-wide halo, 8 rows, the two images of a pair
    if (tc::make_tmap(&tmA, a1, 4, dims, str, box)) return -2;
    uint64_t wd[2] = {64, 288};
    uint64_t ws[1] = {256};
    uint32_t wb[2] = {32, 32};
    if (tc::make_tmap(&tmW, w2, 2, wd, ws, wb, true)) return -3;
  }
  static tc::PerDeviceOnce attr;
  if (attr.need()) {
    if (cudaFuncSetAttribute(conv2_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) return -10;
    attr.mark();
  }
  Params p;
  p.B = B; p.n_tiles = ((B + 1) / 2) * 9; p.training = training; p.keep = keep; p.bias = bias2; p.p2 = p2; p.mask = mask;
  p.st = reinterpret_cast<const mnist::StepState*>(step_state);
  memset(&p.wait, 0, sizeof(p.wait));
  if (wait) p.wait = *wait;
  if (ctas <= 0) ctas = 148;
  if (ctas > p.n_tiles) ctas = p.n_tiles;
  return (int)tc::launch_pdl(conv2_fprop_kernel, dim3(ctas), dim3(THREADS), SMEM_BYTES, (cudaStream_t)stream, tmA, tmW, p);
}

extern "C" int hvd_conv2_fprop(const float* a1, const float* w2, const float* bias2, float* p2, uint8_t* mask,
                               const void* step_state, int B, int training, float keep, int ctas, void* stream) {
  return hvd_conv2_fprop_ex(a1, w2, bias2, p2, mask, step_state, B, training, keep, ctas, stream, nullptr);
}
