// Microbenchmark: how many cycles does one tcgen05.mma.cta_group::1.kind::tf32 of shape M x N x 8 (M = 64 or 128) cost when
// (a) every MMA accumulates into the SAME TMEM columns (a dependent chain, as in an implicit-GEMM K loop) and
// (b) consecutive MMAs rotate over `n_acc` independent accumulators?  The conv kernels' tile shapes (N = 32
// for dgrad, N = 64 for fprop / wgrad) were chosen from this.  Operands are whatever is in shared memory
// (zero-filled): only the issue/complete timing is of interest.
//
// out[blockIdx.x] = cycles between the first issue and the arrival of the final commit.
#include "tc_common.cuh"

namespace {

template <int M, int N>
__global__ void __launch_bounds__(128, 1) mma_probe_kernel(int iters, int n_acc, int a_bytes_step, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                  // 64 KB: [128 rows][128 B] x 4 k-slabs, reused
  uint8_t* sB = smem + 65536;          // 32 KB
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + 65536 + 32768);
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (65536 + 32768) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) { tc::mbar_init(done, 1); tc::fence_barrier_init(); }
  if (warp == 0) tc::tmem_alloc<512>(slot);
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *slot;
  if (warp == 1 && tc::elect_one()) {
    constexpr uint32_t idesc = tc::make_idesc_tf32(M, N, 0, 0);
    const uint64_t dA = tc::desc_kmajor(tc::smem_u32(sA));
    const uint64_t dB = tc::desc_kmajor(tc::smem_u32(sB));
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
      const int acc = i % n_acc;
      // walk over 4 k-slabs x (optionally) different row offsets so the operand fetch is not one hot line
      tc::mma_tf32(tmem + acc * N, tc::desc_advance(dA, (i & 3) * 32 + ((i >> 2) & 3) * a_bytes_step),
                   tc::desc_advance(dB, (i & 3) * 32), idesc, i >= n_acc);
    }
    tc::mma_commit(done);
    tc::mbar_wait(done, 0);
    out[blockIdx.x] = clock64() - t0;
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem);
}

}  // namespace

extern "C" int hvd_mma_probe(int MN, int iters, int n_acc, int a_bytes_step, int ctas, long long* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  const int smem = 65536 + 32768 + 1024 + 256;
  auto launch = [&](auto kern) -> int {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return -1;
    kern<<<ctas, 128, smem, stream>>>(iters, n_acc, a_bytes_step, out);
    return cudaGetLastError() == cudaSuccess ? 0 : -2;
  };
  const int M = MN >> 16, N = MN & 0xffff;          // MN = (M << 16) | N
  if (n_acc < 1 || n_acc * N > 512) return -3;
#define HVD_PROBE_CASE(m, n) if (M == m && N == n) return launch(mma_probe_kernel<m, n>);
  HVD_PROBE_CASE(128, 16) HVD_PROBE_CASE(128, 32) HVD_PROBE_CASE(128, 64) HVD_PROBE_CASE(128, 96)
  HVD_PROBE_CASE(128, 128) HVD_PROBE_CASE(128, 192) HVD_PROBE_CASE(128, 256)
  HVD_PROBE_CASE(64, 32) HVD_PROBE_CASE(64, 64) HVD_PROBE_CASE(64, 96) HVD_PROBE_CASE(64, 128)
  HVD_PROBE_CASE(64, 192) HVD_PROBE_CASE(64, 256)
#undef HVD_PROBE_CASE
  return -4;
}
