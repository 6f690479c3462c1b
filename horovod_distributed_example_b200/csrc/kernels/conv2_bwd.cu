// conv2 backward on tcgen05 / TMEM / TMA (reference layer: Conv2D(64,(3,3)) at mnist_keras.py:75).
//
// dgrad  da1[b,y,x,ci] = relu'(a1) * sum_{kh,kw,co} dY2[b,y-kh,x-kw,co] * W2[kh,kw,ci,co]
//        A 128 x N x 8 TF32 MMA costs ~77 cycles for every N <= 128 (profiles/mma_probe.json), so the
//        "one MMA per tap with N = 32 ci" form wastes 3/4 of the tensor core.  Instead the three kw taps
//        ride in N and only kh is a row shift of A:
//            Z[p][(kw,ci)] = sum_{kh,co} dY2row[p - 26*kh][co] * W2[kh][kw][ci][co]      N = 96, K = 3 x 64
//            da1[p][ci]    = Z[p][0,ci] + Z[p-1][1,ci] + Z[p-2][2,ci]                     (epilogue)
//        p runs over a 26-wide raster of the 26x26 output; dY2 columns 24,25 and rows outside 0..23 are
//        TMA out-of-bounds zero fill, so the x-1 / x-2 neighbours at a row start read zeros and no
//        predicates are needed.  A = one haloed 4-D TMA box ({32 co, 26 w, 8 h}) per channel half; the
//        three kh are row-shifted descriptors into it (start at any row: the tensor core swizzles on
//        absolute smem address bits).  M tiles are 128 consecutive raster positions with a stride of
//        126, so the two neighbours a tile's first outputs need are that tile's own lanes 0 and 1.
//        24 MMAs per 126 outputs instead of 72 per 128 (of which a third were tile padding).
//        B = W2 as [(kw,ci)][co] per (kh, half) (co contiguous = K-major), whole kernel resident in smem.
//        Epilogue: kw reduction by warp shuffles (+ a 3-vector exchange across warp boundaries), conv1's
//        ReLU mask, optional da1 store, and the staging tiles of the fused conv1 weight gradient.
//
// wgrad  dW2[kh,kw,ci,co] = sum_{b,oy,ox} a1[b,oy+kh,ox+kw,ci] * dY2[b,oy,ox,co]
//        the reduction runs over pixels, so BOTH operands are MN-major: A[(kw,ci)][pixel] and
//        B[(kh,co)][pixel] are the activation tensors themselves ([pixel][channel] rows).  The M atoms are
//        the kw shifts of one a1 row (LBO = one row = 128 B, i.e. overlapping atoms) and the N atoms are
//        the three dY2 rows that pair with it, both channel halves (N = 192): one MMA per a1 row does all
//        nine taps.  The bias gradient (column sums of dY2) is summed from the smem tile by the otherwise
//        idle epilogue warps while the tensor core works.  Details at namespace wg.
//        Each CTA reduces its share of the 8-column image strips (dealt round-robin: split-K) and adds its
//        288x64 result into the arena with red.add.v4.
//
// Both are device functions (dg::body, wg::body): each has its own kernel, and conv2_bwd_kernel runs them back to
// back in one launch with an unequal strip/tile split (hvd_conv2_bwd).
#include "mnist_common.cuh"
#include "tc_common.cuh"

namespace {
using namespace mnist;

// Phase timers of CTA 0 (clock64 deltas, one tile in the middle of its range): a handful of clock reads by
// three threads of one CTA, read back with hvd_conv2_bwd_debug_read() by bench/conv_bwd_timing.py.
__device__ long long g_bwd_dbg[64];
#define HVD_DBG_T(slot) do { if (dbg_on) g_bwd_dbg[slot] = clock64(); } while (0)

// ============================================================================ dgrad
namespace dg {
constexpr int STAGES = 3;
constexpr int BOX_W = 26, BOX_H = 8;        // haloed dY2 box: 26 columns (24 + 2 zero), 8 rows (6 + 2 halo)
constexpr int A_STAGE = BOX_W * BOX_H * 128;   // 208 rows x 128 B = 26,624 B (a multiple of 1024)
constexpr int A_TX = A_STAGE;
constexpr int W_BYTES = 18 * 4096;
constexpr int STG_BYTES = 32 * 1024;        // fused conv1-wgrad staging: im2col(x) tile 16 KB + da1 tile 16 KB
constexpr int XCH_BYTES = 2 * 2 * 4 * 96 * 4;   // kw-reduction exchange across warp boundaries: 2 warp sets, double-buffered
constexpr int SMEM_BYTES = W_BYTES + STAGES * A_STAGE + 2 * STG_BYTES + XCH_BYTES + 1024 /*base align*/ + 1024 /*barriers*/ + 256;
constexpr int TMEM_COLS = 256;              // 2 dgrad stages x 96 columns + 32 columns for the conv1 wgrad accumulator
constexpr int ACC_N = 96;                   // (kw, ci)
constexpr int W1_COL = 2 * ACC_N;
constexpr int TILE_STRIDE = 126;            // raster positions owned by a tile (lanes 0,1 only donate, except in tile 0)
constexpr int TILES_PER_IMG = 6;            // 5*126 + 128 = 758 >= 676
constexpr int NPOS = H1 * H1;               // 676

struct Params {
  int B, n_tiles;
  const float* a1;
  float* da1;                 // optional: written only when non-null (the engine does not need it)
  const float* x;             // [B,28,28] network input, for the fused conv1 weight gradient (null = off)
  float* g_w1;                // [9][32]  accumulated with red.add
  float* g_b1;                // [32]
  int n_w, t1;                // tile ownership (see body): n_w < 0 = round-robin over the grid
};

// Row r of an MN-major staging tile (128 B per row, SWIZZLE_128B_BASE32B): 32-byte chunk c lives at
// chunk position c ^ (r & 3).
__device__ __forceinline__ float* stg_chunk(uint8_t* tile, int r, int c) {
  return reinterpret_cast<float*>(tile + r * 128 + ((c ^ (r & 3)) << 5));
}

// The whole dgrad as a device function so that it can run as its own kernel or as the second phase of the
// fused conv2 backward kernel.  `smem` is 1024-byte aligned, `tmem` an allocation of TMEM_COLS columns; every
// thread of the CTA calls it.  It initialises its own mbarriers and ends with a __syncthreads().
__device__ __forceinline__ void body(uint8_t* smem, const uint32_t tmem, const CUtensorMap* tmGp, const CUtensorMap* tmWp,
                                     const Params& p) {
  const CUtensorMap& tmG = *tmGp;
  const CUtensorMap& tmW = *tmWp;
  uint8_t* sW = smem;
  uint8_t* sA = smem + W_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sA + STAGES * A_STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* wfull = bars + 2 * STAGES;
  uint64_t* accf = wfull + 1;
  uint64_t* acce = accf + 2;
  uint64_t* stgf = acce + 2;           // [2] staging tile written by the epilogue
  uint64_t* stge = stgf + 2;           // [2] staging tile consumed by the tensor core
  uint64_t* w1done = stge + 2;
  uint8_t* sStg = sA + STAGES * A_STAGE + 1024;  // the barrier block owns the 1 KB in between (keeps 1024 B alignment)
  float* sXch = reinterpret_cast<float*>(sStg + 2 * STG_BYTES);
  const bool fuse_w1 = p.x != nullptr;
  // Which tiles this CTA owns: t_first + k * t_stride for k < t_count.  Stand-alone: round-robin over the grid.
  // Fused kernel: the first n_w CTAs spent the first phase on wgrad strips and take only t1 tiles, the rest
  // split the remainder evenly (contiguous ranges).
  int t_first, t_stride, t_count;
  {
    const int c = (int)blockIdx.x, G = (int)gridDim.x;
    if (p.n_w < 0) {
      t_first = c; t_stride = G;
      t_count = c < p.n_tiles ? (p.n_tiles - c + G - 1) / G : 0;
    } else {
      t_stride = 1;
      if (c < p.n_w) { t_first = c * p.t1; t_count = p.t1; }
      else {
        const int rest = p.n_tiles - p.n_w * p.t1, pure = G - p.n_w, j = c - p.n_w;
        const int t2 = rest / pure, rem = rest - t2 * pure;
        t_first = p.n_w * p.t1 + j * t2 + (j < rem ? j : rem);
        t_count = t2 + (j < rem ? 1 : 0);
      }
    }
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) { tc::tma_prefetch_desc(&tmG); tc::tma_prefetch_desc(&tmW); }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    tc::mbar_init(wfull, 1);
    for (int s = 0; s < 2; s++) {
      tc::mbar_init(&accf[s], 1); tc::mbar_init(&acce[s], 128);
      tc::mbar_init(&stgf[s], 128); tc::mbar_init(&stge[s], 1);
    }
    tc::mbar_init(w1done, 1);
    tc::fence_barrier_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  if (warp == 0) {
    if (tc::elect_one()) {
      tc::mbar_expect_tx(wfull, W_BYTES);
      // smem W layout: block (kh, half, kw) of [32 ci][32 co] -> for a fixed (kh, half) the 96 rows (kw, ci)
      // are contiguous = one N = 96 K-major operand
      for (int t = 0; t < 9; t++)
        for (int h = 0; h < 2; h++)
          tc::tma_load_2d(sW + (((t / 3) * 2 + h) * 3 + (t % 3)) * 4096, &tmW, wfull, 32 * h, t * 32);
      int it = 0;
      for (int kt = 0; kt < t_count; kt++) {
        const int tile = t_first + kt * t_stride;
        const int b = tile / TILES_PER_IMG, ybase = ((tile % TILES_PER_IMG) * TILE_STRIDE) / H1;
        for (int half = 0; half < 2; half++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
          tc::mbar_expect_tx(&full[s], A_TX);
          // haloed dY2 rows ybase-2 .. ybase+5, columns 0..25: everything outside 24x24 is TMA zero fill
          // (= the "full" convolution's padding and the raster's two pad columns)
          tc::tma_load_4d(sA + s * A_STAGE, &tmG, &full[s], 32 * half, 0, ybase - 2, b);
        }
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(128, ACC_N, 0, 0);
      constexpr uint32_t idesc_w1 = tc::make_idesc_tf32(128, 32, 1, 1);
      const uint64_t dW0 = tc::desc_kmajor(tc::smem_u32(sW));
      // Fused conv1 weight gradient: dW1[tap][c] (+ db1 as "tap" 9) = sum over the tile's 128 pixels of
      // im2col(x)[pixel][tap] * da1[pixel][c].  Both staging tiles are [pixel][32] rows written by the
      // epilogue = MN-major operands with K = pixels: 16 MMAs (K=8) per tile into TMEM columns W1_COL..+31.
      auto w1_mma = [&](int tj) {
        const int sb = tj & 1;
        tc::mbar_wait(&stgf[sb], (tj >> 1) & 1);
        tc::tc_fence_after();
        const uint32_t st = tc::smem_u32(sStg + sb * STG_BYTES);
        const uint64_t dA = tc::make_smem_desc(st, 0, 512, 1);            // LBO 0: the four M atoms coincide
        const uint64_t dB = tc::make_smem_desc(st + 16384, 0, 512, 1);
#pragma unroll
        for (int ks = 0; ks < 16; ks++)
          tc::mma_tf32(tmem + W1_COL, tc::desc_advance(dA, ks * 1024), tc::desc_advance(dB, ks * 1024), idesc_w1, (tj | ks) != 0);
        tc::mma_commit(&stge[sb]);
      };
      tc::mbar_wait(wfull, 0);
      int it = 0, ti = 0;
      for (; ti < t_count; ti++) {
        const int tile = t_first + ti * t_stride;
        const int as = ti & 1;
        const bool dbg_on = blockIdx.x == 0 && (ti == 2 || ti == 3);
        HVD_DBG_T(8 + (ti & 1) * 8);
        tc::mbar_wait(&acce[as], ((ti >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        HVD_DBG_T(9 + (ti & 1) * 8);
        const int p0 = (tile % TILES_PER_IMG) * TILE_STRIDE;
        const int row0 = p0 - (p0 / H1) * H1;            // raster offset of lane 0 inside the box (box row 2 = ybase)
        for (int half = 0; half < 2; half++, it++) {
          const int s = it % STAGES;
          tc::mbar_wait(&full[s], (it / STAGES) & 1);
          tc::tc_fence_after();
          HVD_DBG_T(12 + (ti & 1) * 8 + half);
          const uint64_t da0 = tc::make_smem_desc(tc::smem_u32(sA + s * A_STAGE), 16, 1024, 2);
#pragma unroll
          for (int kh = 0; kh < 3; kh++) {
            // output raster position p needs dY2 row (y - kh) = box raster p - p0 + row0 + (2 - kh) * 26
            const uint64_t da = tc::desc_advance(da0, (row0 + (2 - kh) * BOX_W) * 128);
            const uint64_t dw = tc::desc_advance(dW0, (kh * 2 + half) * 3 * 4096);
#pragma unroll
            for (int k = 0; k < 4; k++)
              tc::mma_tf32(tmem + as * ACC_N, tc::desc_advance(da, k * 32), tc::desc_advance(dw, k * 32), idesc,
                           (half | kh | k) != 0);
          }
          tc::mma_commit(&empty[s]);
        }
        tc::mma_commit(&accf[as]);
        HVD_DBG_T(10 + (ti & 1) * 8);
        // conv1 wgrad of the tile TWO back: its epilogue (the other tile of the same warp set) has finished
        // staging by the time this tile's main MMAs are queued, so the issue thread never waits here long
        if (fuse_w1 && ti > 1) w1_mma(ti - 2);
        HVD_DBG_T(11 + (ti & 1) * 8);
      }
      if (fuse_w1) {
        for (int tj = ti > 2 ? ti - 2 : 0; tj < ti; tj++) w1_mma(tj);
        tc::mma_commit(w1done);
      }
    }
  } else if (warp >= 4) {
    // Two epilogue warp sets (warps 4-7 and 8-11) alternate tiles: set s owns accumulator stage s, staging
    // buffer s and its own exchange buffers, so one tile's long-latency epilogue (TMEM loads, a1 / x loads
    // from L2, staging stores) overlaps the other's.
    const int q = warp & 3;
    const int set = (warp - 4) >> 2;
    const int r = q * 32 + lane;               // accumulator row = raster position p0 + r
    int ti = set;
    for (; ti < t_count; ti += 2) {
      const int tile = t_first + ti * t_stride;
      const int as = ti & 1;
      const bool dbg_on = blockIdx.x == 0 && threadIdx.x == 128 && ti == 2;
      const int b = tile / TILES_PER_IMG, tt = tile % TILES_PER_IMG;
      const int pos = tt * TILE_STRIDE + r;
      const int y = pos / H1, x = pos - y * H1;
      // lanes 0,1 of tiles 1..5 repeat the previous tile's last two positions: donors only
      const bool valid = pos < NPOS && (r >= 2 || tt == 0);
      const long long off = ((long long)b * NPOS + pos) * C1;
      // Everything this tile needs from L2 is requested BEFORE waiting for the accumulator: conv1's
      // activations (reduced to a 32-bit ReLU mask as soon as they land) and the 3x3 input patch of the
      // fused conv1 weight gradient.
      uint32_t relu = 0;
      float xin[10];
#pragma unroll
      for (int k = 0; k < 10; k++) xin[k] = 0.f;
      {
        float4 act[8];
        if (valid) {
          const float4* ap = reinterpret_cast<const float4*>(p.a1 + off);
#pragma unroll
          for (int j = 0; j < 8; j++) act[j] = __ldg(ap + j);
          if (fuse_w1) {
            const float* xi = p.x + ((long long)b * IMG + y) * IMG + x;
#pragma unroll
            for (int kh = 0; kh < 3; kh++)
#pragma unroll
              for (int kw = 0; kw < 3; kw++) xin[kh * 3 + kw] = __ldg(xi + kh * IMG + kw);
            xin[9] = 1.0f;                                  // the bias "tap"
          }
        }
        HVD_DBG_T(0);
        tc::mbar_wait(&accf[as], (ti >> 1) & 1);
        tc::tc_fence_after();
        HVD_DBG_T(1);
        if (valid) {
#pragma unroll
          for (int j = 0; j < 8; j++)
            relu |= ((act[j].x > 0.f ? 1u : 0u) | (act[j].y > 0.f ? 2u : 0u) | (act[j].z > 0.f ? 4u : 0u) | (act[j].w > 0.f ? 8u : 0u)) << (4 * j);
        }
      }
      // da1[p] = Z[p][kw=0] + Z[p-1][kw=1] + Z[p-2][kw=2]: neighbours are the lanes below (shuffles); the
      // first two lanes of warps 1..3 take them from the previous warp through a small smem exchange
      float v[32], t1[32], t2[32];
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16) + as * ACC_N;
      const int xb = set * 2 + ((ti >> 1) & 1);
      float* xw = sXch + (xb * 4 + q) * 96;
      tc::tmem_ld_x32(trow, v);
      tc::tmem_ld_x32(trow + 32, t1);
      tc::tmem_ld_x32(trow + 64, t2);
      tc::tmem_ld_wait();
      HVD_DBG_T(6);
      tc::tc_fence_before();
      tc::mbar_arrive(&acce[as]);                           // the accumulator stage is free for tile ti+2
      if (lane >= 30) {
        float4* d = reinterpret_cast<float4*>(xw + (lane == 30 ? 32 : 64));
#pragma unroll
        for (int j = 0; j < 8; j++) d[j] = make_float4(t2[4 * j], t2[4 * j + 1], t2[4 * j + 2], t2[4 * j + 3]);
        if (lane == 31) {
#pragma unroll
          for (int j = 0; j < 8; j++) reinterpret_cast<float4*>(xw)[j] = make_float4(t1[4 * j], t1[4 * j + 1], t1[4 * j + 2], t1[4 * j + 3]);
        }
      }
#pragma unroll
      for (int j = 0; j < 32; j++) {
        const float u1 = __shfl_up_sync(0xffffffffu, t1[j], 1);
        const float u2 = __shfl_up_sync(0xffffffffu, t2[j], 2);
        v[j] += (lane >= 1 ? u1 : 0.f) + (lane >= 2 ? u2 : 0.f);
      }
      HVD_DBG_T(7);
      if (set == 0) asm volatile("bar.sync 1, 128;" ::: "memory");      // the four warps of this set
      else          asm volatile("bar.sync 2, 128;" ::: "memory");
      if (q >= 1 && lane < 2) {
        const float4* xr = reinterpret_cast<const float4*>(sXch + (xb * 4 + q - 1) * 96);
#pragma unroll
        for (int j = 0; j < 8; j++) {
          float4 a;
          if (lane == 0) {
            const float4 a1v = xr[j], a2v = xr[8 + j];
            a = make_float4(a1v.x + a2v.x, a1v.y + a2v.y, a1v.z + a2v.z, a1v.w + a2v.w);
          } else {
            a = xr[16 + j];
          }
          v[4 * j] += a.x; v[4 * j + 1] += a.y; v[4 * j + 2] += a.z; v[4 * j + 3] += a.w;
        }
      }
      HVD_DBG_T(2);
#pragma unroll
      for (int j = 0; j < 32; j++) v[j] = ((relu >> j) & 1u) ? v[j] : 0.f;     // conv1's ReLU backward (0 for invalid lanes)
      if (valid && p.da1) {
        float4* dst = reinterpret_cast<float4*>(p.da1 + off);
#pragma unroll
        for (int j = 0; j < 8; j++) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
      HVD_DBG_T(3);
      if (fuse_w1) {
        const int sb = ti & 1;
        tc::mbar_wait(&stge[sb], ((ti >> 1) & 1) ^ 1);    // the tensor core is done with this staging tile
        HVD_DBG_T(4);
        uint8_t* tA = sStg + sb * STG_BYTES;
        uint8_t* tB = tA + 16384;
        float4* a0 = reinterpret_cast<float4*>(stg_chunk(tA, r, 0));
        a0[0] = make_float4(xin[0], xin[1], xin[2], xin[3]);
        a0[1] = make_float4(xin[4], xin[5], xin[6], xin[7]);
        float4* a1c = reinterpret_cast<float4*>(stg_chunk(tA, r, 1));
        a1c[0] = make_float4(xin[8], xin[9], 0.f, 0.f);
        a1c[1] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int c = 2; c < 4; c++) {
          float4* z = reinterpret_cast<float4*>(stg_chunk(tA, r, c));
          z[0] = make_float4(0.f, 0.f, 0.f, 0.f);
          z[1] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int c = 0; c < 4; c++) {
          float4* d = reinterpret_cast<float4*>(stg_chunk(tB, r, c));
          d[0] = make_float4(v[8 * c], v[8 * c + 1], v[8 * c + 2], v[8 * c + 3]);      // masked: 0 on invalid lanes
          d[1] = make_float4(v[8 * c + 4], v[8 * c + 5], v[8 * c + 6], v[8 * c + 7]);
        }
        tc::fence_proxy_async();                          // generic-proxy smem writes -> visible to the tensor core
        tc::mbar_arrive(&stgf[sb]);
      }
      HVD_DBG_T(5);
    }
    if (fuse_w1 && warp == 4) {
      // rows 0..8 of the conv1-wgrad accumulator are the 9 taps, row 9 is the bias gradient
      tc::mbar_wait(w1done, 0);
      tc::tc_fence_after();
      float w[32];
      tc::tmem_ld_x32(tmem + W1_COL, w);
      tc::tmem_ld_wait();
      if (t_count > 0 && lane < 10) {
        float* d = lane < 9 ? p.g_w1 + lane * 32 : p.g_b1;
#pragma unroll
        for (int j = 0; j < 32; j += 4) tc::red_add_f4(d + j, w[j], w[j + 1], w[j + 2], w[j + 3]);
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
}

__global__ void __launch_bounds__(384, 1)
conv2_dgrad_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmW, Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5;
  if (warp == 2) tc::tmem_alloc<TMEM_COLS>(&tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  body(smem, tmem, &tmG, &tmW, p);
  if (warp == 2) tc::tmem_dealloc<TMEM_COLS>(tmem);
}
}  // namespace dg

// ============================================================================ wgrad
namespace wg {
// One work item = an 8-column strip of one image over its full height:
//   A: a1 halo box {32 ci, 10 w, 26 h}            -> raster [h][w], row (h, w) at (h*10 + w) * 128 B
//   B: dY2 box {32 co, 8 w, 2 halves, 28 h} from a 5-D view of dY2 whose third dimension is the channel
//      half (stride 128 B) -> raster [h][half][w]: the 8 pixels of (row h, half) are one contiguous 1 KB
//      K-slab at ((h*2 + half) * 8) * 128 B.  Rows -2,-1 and 24,25 are TMA zero fill.
// One MMA per a1 row hr (K = its 8 pixels): M atoms = the kw shifts of that row (LBO = one row, the 4th atom
// is a dummy), N atoms = (kh', half) for the dY2 rows hr-2+kh' (LBO = 1 KB):
//   D[(kw,ci)][(kh',half,co)] += a1[hr][x+kw][ci] * dY2[hr-2+kh'][x][co]      kh = 2 - kh'
// 26 MMAs of N = 192 (~97 cycles each) per strip, where one MMA per (row, kh) of N = 64 cost 3 x 77.
constexpr int HALO_ROWS = 26 * 10;
constexpr int HALO_TX = HALO_ROWS * 128;    // 33,280 B
constexpr int HALO_BYTES = 33 * 1024;       // + 4 slack rows: the dummy 4th atom of the last row reads one row past the box
constexpr int DY_ROWS = 28 * 2 * 8;
constexpr int DY_BYTES = DY_ROWS * 128;     // 57,344 B
constexpr int STAGE = HALO_BYTES + DY_BYTES;
constexpr int STAGE_TX = HALO_TX + DY_BYTES;
constexpr int STAGES = 2;
constexpr int SMEM_BYTES = STAGES * STAGE + 1024 /*base align*/ + 1024 /*barriers*/ + 256;
constexpr int ACC_N = 192;
constexpr int TMEM_COLS = 256;

struct Params {
  int B, n_strips, n_w;
  float* g_w;                               // [288][64]
  float* g_b;                               // [64]
  const StepState* st;                      // optional: publish the step's (loss, accuracy, step) ...
  float* host_ring;                         // ... into this page-locked host ring [16][4]
};

// Device-function form (own kernel, or first phase of the fused conv2 backward kernel): see dg::body.
// Uses warps 0..7 only.  Strips are dealt round-robin over the first n_w CTAs: strip = blockIdx.x + i * n_w.
__device__ __forceinline__ void body(uint8_t* smem, const uint32_t tmem, const CUtensorMap* tmAp, const CUtensorMap* tmGp,
                                     const Params& p) {
  const CUtensorMap& tmA = *tmAp;
  const CUtensorMap& tmG = *tmGp;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* accf = bars + 2 * STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nw = p.n_w > 0 ? p.n_w : (int)gridDim.x;      // CTAs that take strips (the fused kernel uses a subset)
  const int ns = (int)blockIdx.x < nw && (int)blockIdx.x < p.n_strips ? (p.n_strips - (int)blockIdx.x + nw - 1) / nw : 0;
  if (warp == 0 && lane == 0) { tc::tma_prefetch_desc(&tmA); tc::tma_prefetch_desc(&tmG); }
  if (warp == 1 && lane == 0) {
    // a stage is free again when the tensor core (1 commit) and the four bias-sum warps (1 arrive each) are done
    for (int s = 0; s < STAGES; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 5); }
    tc::mbar_init(accf, 1);
    tc::fence_barrier_init();
  }
  // the slack rows after each halo tile are read by the dummy 4th atom: keep them finite
  for (int i = threadIdx.x; i < STAGES * 4 * 32; i += blockDim.x)
    reinterpret_cast<float*>(smem + (i / 128) * STAGE + HALO_TX)[i % 128] = 0.f;
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();

  if (warp == 0) {
    if (tc::elect_one()) {
      for (int i = 0; i < ns; i++) {
        const int s = i % STAGES;
        tc::mbar_wait(&empty[s], ((i / STAGES) & 1) ^ 1);
        tc::mbar_expect_tx(&full[s], STAGE_TX);
        const int strip = (int)blockIdx.x + i * nw;
        const int b = strip / 3, px = strip % 3;
        uint8_t* st = smem + s * STAGE;
        tc::tma_load_4d(st, &tmA, &full[s], 0, px * 8, 0, b);                      // a1 columns px*8 .. px*8+9, all 26 rows
        tc::tma_load_5d(st + HALO_BYTES, &tmG, &full[s], 0, px * 8, 0, -2, b);     // dY2 rows -2..25, both halves
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(128, ACC_N, 1, 1);
      for (int i = 0; i < ns; i++) {
        const int s = i % STAGES;
        tc::mbar_wait(&full[s], (i / STAGES) & 1);
        tc::tc_fence_after();
        const uint32_t halo = tc::smem_u32(smem + s * STAGE);
        const uint64_t da0 = tc::make_smem_desc(halo, 128, 512, 1);
        const uint64_t db0 = tc::make_smem_desc(halo + HALO_BYTES, 1024, 512, 1);
#pragma unroll 2
        for (int hr = 0; hr < 26; hr++)
          tc::mma_tf32(tmem, tc::desc_advance(da0, hr * 1280), tc::desc_advance(db0, hr * 2048), idesc, (i | hr) != 0);
        tc::mma_commit(&empty[s]);
      }
      tc::mma_commit(accf);
    }
  } else if (warp == 3) {
    // The head kernel finished long ago (stream order): hand the step's result to the host from here, off
    // the critical path.  ONE aligned 16-byte store = one posted PCIe write, so the host never sees the step
    // number without the values; nothing on the compute stream waits for it.
    if (blockIdx.x == 0 && lane == 0 && p.host_ring != nullptr) {
      const unsigned step = p.st->step;
      *reinterpret_cast<float4*>(p.host_ring + (step & 15u) * 4) =
          make_float4(p.st->last_loss, p.st->last_correct, __uint_as_float(step), 0.f);
    }
  } else if (warp >= 4 && warp < 8 && ns > 0) {
    // ---- while the tensor core works: conv2's bias gradient = column sums of the dY2 tile (every dY2 pixel
    // is in exactly one strip; the halo rows are zeros).  Thread t reads the 16-byte chunk at position t%8 of
    // rows t/8, t/8+16, ...: r & 3 and the channel half (r >> 3) & 1 are the same for all of them, so after
    // undoing the 32-byte-atom swizzle each thread owns four fixed channels.
    const int q = warp & 3;
    const int t = threadIdx.x - 128;
    const int pc = t & 7, rbase = t >> 3;
    float4 bs = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < ns; i++) {
      const int s = i % STAGES;
      tc::mbar_wait(&full[s], (i / STAGES) & 1);
      const uint8_t* dy = smem + s * STAGE + HALO_BYTES;
#pragma unroll 4
      for (int k = 0; k < DY_ROWS / 16; k++) {
        const float4 v = *reinterpret_cast<const float4*>(dy + (rbase + 16 * k) * 128 + pc * 16);
        bs.x += v.x; bs.y += v.y; bs.z += v.z; bs.w += v.w;
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&empty[s]);
    }
    {
      const int half = (rbase >> 3) & 1;
      const int c32 = (pc >> 1) ^ (rbase & 3);
      tc::red_add_f4(p.g_b + half * 32 + c32 * 8 + (pc & 1) * 4, bs.x, bs.y, bs.z, bs.w);
    }
    // ---- epilogue: accumulator row = (kw = q, ci = lane); column block a = kh'*2 + half
    tc::mbar_wait(accf, 0);
    tc::tc_fence_after();
    if (q != 3) {                             // the 4th M atom is the dummy shift (warp-uniform skip)
#pragma unroll 1
      for (int a6 = 0; a6 < 6; a6++) {
        // every CTA adds into the same 288x64 block: start each one at a different column block so the
        // red.adds of CTAs that finish together land on different addresses
        const int a = (a6 + blockIdx.x) % 6;
        float v[32];
        tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + a * 32, v);
        tc::tmem_ld_wait();
        const int kh = 2 - (a >> 1);
        float* d = p.g_w + ((long long)((kh * 3 + q) * 32 + lane)) * 64 + (a & 1) * 32;
#pragma unroll
        for (int j = 0; j < 32; j += 4) tc::red_add_f4(d + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
  }
  tc::fence_proxy_async();      // generic-proxy reads of the stages are ordered before whatever TMA writes there next
  tc::tc_fence_before();
  __syncthreads();
}

__global__ void __launch_bounds__(256, 1)
conv2_wgrad_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmG, Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5;
  if (warp == 2) tc::tmem_alloc<TMEM_COLS>(&tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  body(smem, tmem, &tmA, &tmG, p);
  if (warp == 2) tc::tmem_dealloc<TMEM_COLS>(tmem);
}
}  // namespace wg

// ============================================================================ both in one launch
// Phase 1: the first n_w CTAs reduce the wgrad strips (every CTA that takes strips pays 73 KB of red.adds
// for its 288x64 partial result, so not all of them do).  Phase 2: all CTAs compute dgrad tiles in the same
// shared memory and TMEM - the wgrad CTAs only t1 each, the others the rest.  Compared with two kernels side
// by side on a fixed SM split this balances the work to within one tile per CTA, lets the red.adds drain
// while dgrad computes, and saves a launch and a stream join.
constexpr int FUSED_SMEM = dg::SMEM_BYTES > wg::SMEM_BYTES ? dg::SMEM_BYTES : wg::SMEM_BYTES;
static_assert(dg::TMEM_COLS == wg::TMEM_COLS, "the two phases share one TMEM allocation");

__global__ void __launch_bounds__(384, 1)
conv2_bwd_kernel(const __grid_constant__ CUtensorMap tmA_w, const __grid_constant__ CUtensorMap tmG_w,
                 const __grid_constant__ CUtensorMap tmG_d, const __grid_constant__ CUtensorMap tmW_d, wg::Params pw,
                 dg::Params pd) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5;
  if (warp == 2) tc::tmem_alloc<dg::TMEM_COLS>(&tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  wg::body(smem, tmem, &tmA_w, &tmG_w, pw);
  dg::body(smem, tmem, &tmG_d, &tmW_d, pd);
  if (warp == 2) tc::tmem_dealloc<dg::TMEM_COLS>(tmem);
}

}  // namespace

// da1 may be NULL.  When x is non-NULL the conv1 weight/bias gradients are accumulated (red.add) into
// g_w1 [9][32] / g_b1 [32] inside the same kernel (no separate conv1-wgrad pass, no da1 round trip).
extern "C" int hvd_conv2_dgrad(const float* dy2, const float* w2, const float* a1, float* da1, int B, int ctas,
                               void* stream, const float* x, float* g_w1, float* g_b1) {
  CUtensorMap tmG, tmW;
  {
    uint64_t dims[4] = {64, 24, 24, (uint64_t)B};
    uint64_t str[3] = {256, 24 * 256, 24 * 24 * 256};
    uint32_t box[4] = {32, dg::BOX_W, dg::BOX_H, 1};
    if (tc::make_tmap(&tmG, dy2, 4, dims, str, box)) return -2;
    uint64_t wd[2] = {64, 288};
    uint64_t ws[1] = {256};
    uint32_t wb[2] = {32, 32};
    if (tc::make_tmap(&tmW, w2, 2, wd, ws, wb)) return -3;
  }
  static tc::PerDeviceOnce attr;
  if (attr.need()) {
    if (cudaFuncSetAttribute(dg::conv2_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dg::SMEM_BYTES) != cudaSuccess) return -10;
    attr.mark();
  }
  dg::Params p{B, B * dg::TILES_PER_IMG, a1, da1, x, g_w1, g_b1, -1, 0};
  if (ctas <= 0) ctas = 148;
  if (ctas > p.n_tiles) ctas = p.n_tiles;
  return (int)tc::launch_pdl(dg::conv2_dgrad_kernel, dim3(ctas), dim3(384), dg::SMEM_BYTES, (cudaStream_t)stream, tmG, tmW, p);
}

// g_w [3][3][32][64] and g_b [64] are ACCUMULATED into (red.add): the caller zeroes them before the step.
extern "C" int hvd_conv2_bwd_debug_read(long long* out64) {
  return (int)cudaMemcpyFromSymbol(out64, g_bwd_dbg, sizeof(long long) * 64);
}

extern "C" int hvd_conv2_wgrad_ex(const float* a1, const float* dy2, float* g_w, float* g_b, int B, int ctas, void* stream,
                                  const void* step_state, float* host_ring) {
  CUtensorMap tmA, tmG;
  {
    uint64_t dims[4] = {32, 26, 26, (uint64_t)B};
    uint64_t str[3] = {128, 26 * 128, 26 * 26 * 128};
    uint32_t hbox[4] = {32, 10, 26, 1};
    if (tc::make_tmap(&tmA, a1, 4, dims, str, hbox, true)) return -2;
    // dY2 [B][24][24][64] seen as {32 co, 24 w, 2 halves, 24 h, B}: the channel half is its own dimension so
    // one box lands in smem as [h][half][w]
    uint64_t gd[5] = {32, 24, 2, 24, (uint64_t)B};
    uint64_t gs[4] = {256, 128, 24 * 256, 24 * 24 * 256};
    uint32_t box[5] = {32, 8, 2, 28, 1};
    if (tc::make_tmap(&tmG, dy2, 5, gd, gs, box, true)) return -3;
  }
  static tc::PerDeviceOnce attr;
  if (attr.need()) {
    if (cudaFuncSetAttribute(wg::conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wg::SMEM_BYTES) != cudaSuccess) return -10;
    attr.mark();
  }
  wg::Params p;
  p.B = B; p.n_strips = B * 3; p.n_w = 0;
  if (ctas <= 0) ctas = 128;
  if (ctas > p.n_strips) ctas = p.n_strips;
  p.g_w = g_w; p.g_b = g_b;
  p.st = reinterpret_cast<const mnist::StepState*>(step_state);
  p.host_ring = step_state ? host_ring : nullptr;
  return (int)tc::launch_pdl(wg::conv2_wgrad_kernel, dim3(ctas), dim3(256), wg::SMEM_BYTES, (cudaStream_t)stream, tmA, tmG, p);
}

// conv2 wgrad + dgrad (+ fused conv1 wgrad) in ONE launch: g_w2 [3][3][32][64], g_b2 [64], g_w1 [9][32], g_b1 [32]
// are accumulated (zero them first); step_state / host_ring as in hvd_conv2_wgrad_ex (may be NULL).
extern "C" int hvd_conv2_bwd(const float* dy2, const float* w2, const float* a1, const float* x, float* g_w2, float* g_b2,
                             float* g_w1, float* g_b1, int B, int ctas, void* stream, const void* step_state,
                             float* host_ring, int wgrad_ctas, int wgrad_cta_tiles) {
  CUtensorMap tmA_w, tmG_w, tmG_d, tmW_d;
  {
    uint64_t ad[4] = {32, 26, 26, (uint64_t)B};
    uint64_t as[3] = {128, 26 * 128, 26 * 26 * 128};
    uint32_t hbox[4] = {32, 10, 26, 1};
    if (tc::make_tmap(&tmA_w, a1, 4, ad, as, hbox, true)) return -2;
    uint64_t gd5[5] = {32, 24, 2, 24, (uint64_t)B};
    uint64_t gs5[4] = {256, 128, 24 * 256, 24 * 24 * 256};
    uint32_t box5[5] = {32, 8, 2, 28, 1};
    if (tc::make_tmap(&tmG_w, dy2, 5, gd5, gs5, box5, true)) return -3;
    uint64_t gd[4] = {64, 24, 24, (uint64_t)B};
    uint64_t gs[3] = {256, 24 * 256, 24 * 24 * 256};
    uint32_t box[4] = {32, dg::BOX_W, dg::BOX_H, 1};
    if (tc::make_tmap(&tmG_d, dy2, 4, gd, gs, box)) return -4;
    uint64_t wd[2] = {64, 288};
    uint64_t ws[1] = {256};
    uint32_t wb[2] = {32, 32};
    if (tc::make_tmap(&tmW_d, w2, 2, wd, ws, wb)) return -5;
  }
  static tc::PerDeviceOnce attr;
  if (attr.need()) {
    if (cudaFuncSetAttribute(conv2_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FUSED_SMEM) != cudaSuccess) return -10;
    attr.mark();
  }
  if (ctas <= 0) ctas = 148;
  const int n_tiles = B * dg::TILES_PER_IMG;
  // the first n_w CTAs reduce the wgrad strips (a strip costs about 0.6 dgrad tiles) and then take t1 tiles each
  int n_w = wgrad_ctas > 0 ? wgrad_ctas : (ctas * 48) / 148;
  if (n_w < 1) n_w = 1;
  if (n_w > ctas - 1) n_w = ctas > 1 ? ctas - 1 : 1;
  int t1 = wgrad_cta_tiles >= 0 ? wgrad_cta_tiles : 2;
  if (n_w * t1 > n_tiles) t1 = n_tiles / n_w;
  if (ctas == 1) { n_w = 1; t1 = n_tiles; }
  wg::Params pw;
  pw.B = B; pw.n_strips = B * 3; pw.n_w = n_w; pw.g_w = g_w2; pw.g_b = g_b2;
  pw.st = reinterpret_cast<const mnist::StepState*>(step_state);
  pw.host_ring = step_state ? host_ring : nullptr;
  dg::Params pd{B, n_tiles, a1, nullptr, x, g_w1, g_b1, n_w, t1};
  return (int)tc::launch_pdl(conv2_bwd_kernel, dim3(ctas), dim3(384), FUSED_SMEM, (cudaStream_t)stream, tmA_w, tmG_w, tmG_d, tmW_d, pw, pd);
}

extern "C" int hvd_conv2_wgrad(const float* a1, const float* dy2, float* g_w, float* g_b, int B, int ctas, void* stream) {
  return hvd_conv2_wgrad_ex(a1, dy2, g_w, g_b, B, ctas, stream, nullptr, nullptr);
}
