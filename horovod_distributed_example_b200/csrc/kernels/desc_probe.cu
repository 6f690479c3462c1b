// Bring-up probe: does a UMMA shared-memory descriptor whose start address is shifted by R rows
// (R not a multiple of the swizzle period) and whose group stride (SBO) is not a multiple of the
// swizzle period still address a TMA-written swizzled tile correctly, and with which `base_offset`?
// The halo-reuse convolution kernels depend on the answer (one haloed tile serves all 9 taps).
//
// One CTA.  A tile: `rows_a` rows of 128 B loaded by one TMA box; the MMA reads M=128 (K-major: rows
// are M; 16 groups of 8 rows, group g at start + g*sbo) or K=32 (MN-major: rows are K; 8 groups of
// 4 rows).  B is a plain K-major [64][32] tile.  D[128][64] is stored for comparison on the host.
#include "tc_common.cuh"

namespace {

struct ProbeParams {
  int a_mn;            // 0: A K-major (rows = M), 1: A MN-major (rows = K)
  int rows_a;          // rows in the loaded A tile
  int shift_rows;      // descriptor start = tile + shift_rows*128
  uint32_t lbo, sbo;
  int base_offset;     // value for descriptor bits [49,52)
  float* D;            // [128][64]
};

__global__ void __launch_bounds__(128, 1)
desc_probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, ProbeParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                  // up to 64 KB
  uint8_t* sB = smem + 65536;          // 8 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 65536 + 8192);
  uint64_t* done = bar + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::mbar_init(done, 1); tc::fence_barrier_init(); }
  if (warp == 0) tc::tmem_alloc<64>(slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *slot;
  if (warp == 1 && tc::elect_one()) {
    tc::mbar_expect_tx(bar, p.rows_a * 128 + 8192);
    tc::tma_load_2d(sA, &tmA, bar, 0, 0);
    tc::tma_load_2d(sB, &tmB, bar, 0, 0);
    tc::mbar_wait(bar, 0);
    tc::tc_fence_after();
    const uint32_t idesc = tc::make_idesc_tf32(128, 64, p.a_mn, 0);
    const uint32_t a0 = tc::smem_u32(sA) + p.shift_rows * 128, b0 = tc::smem_u32(sB);
    for (int k = 0; k < 4; k++) {
      uint64_t da = p.a_mn ? tc::make_smem_desc(a0 + k * 8 * 128, p.lbo, p.sbo, 1)      // 8 k-rows per MMA
                           : tc::make_smem_desc(a0 + k * 32, p.lbo, p.sbo, 2);
      da |= (uint64_t)(p.base_offset & 7) << 49;
      tc::mma_tf32(tmem, da, tc::desc_kmajor(b0 + k * 32), idesc, k != 0);
    }
    tc::mma_commit(done);
  }
  tc::mbar_wait(done, 0);
  tc::tc_fence_after();
  for (int c = 0; c < 64; c += 32) {
    float v[32];
    tc::tmem_ld_x32(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    tc::tmem_ld_wait();
    for (int j = 0; j < 32; j++) p.D[(warp * 32 + lane) * 64 + c + j] = v[j];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<64>(tmem);
}

}  // namespace

// Ag: [rows_a][32] fp32 (row = 128 B).  Bg: [64][32] fp32 (K-major B).  D: [128][64].
extern "C" int hvd_desc_probe(const float* Ag, const float* Bg, float* D, int a_mn, int rows_a, int shift_rows,
                              unsigned lbo, unsigned sbo, int base_offset, void* stream) {
  CUtensorMap ta, tb;
  uint64_t dims[2] = {32, (uint64_t)rows_a};
  uint64_t str[1] = {128};
  uint32_t box[2] = {32, (uint32_t)rows_a};
  if (rows_a > 256 || tc::make_tmap(&ta, Ag, 2, dims, str, box, a_mn != 0)) return -2;
  uint64_t bd[2] = {32, 64};
  uint32_t bb[2] = {32, 64};
  if (tc::make_tmap(&tb, Bg, 2, bd, str, bb, false)) return -3;
  const int smem = 65536 + 8192 + 1024 + 128;
  static tc::PerDeviceOnce attr;
  if (attr.need()) {
    if (cudaFuncSetAttribute(desc_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -10;
    attr.mark();
  }
  ProbeParams p{a_mn, rows_a, shift_rows, lbo, sbo, base_offset, D};
  desc_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(ta, tb, p);
  return (int)cudaGetLastError();
}
