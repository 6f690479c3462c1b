// Descriptors of the pipelined data-parallel step: the gradient of the big
// Dense kernel never sits in a local gradient buffer waiting for a collective.
//
//   1. The GEMM that produces it pushes every output tile straight into the receive area of the rank that
//      owns those rows (tc_gemm EPI_RS_PUSH: GEMM -> reduce-scatter in one kernel, over peer memory), while
//      the conv backward is still running.
//   2. The step's tail kernel (pipeline_tail.cu) reduces + updates the owned rows locally and runs a
//      barrier-free LL one-shot allreduce + update of the small tensors - the only cross-GPU latency left on
//      the critical path.
//   3. The all-gather of the updated rows is issued at the START of the next step by a few small CTAs that
//      share the SMs with the conv forward kernels (hvd_allgather_kernel); each stamps a version flag on
//      every peer, and the first kernel of that step that reads the big weights (tc_gemm wait hook) waits for
//      the flags instead of anybody sitting in an exit barrier.
//
// Capability parity: hvd.DistributedOptimizer (reference mnist_keras.py:87, tensorflow2_keras_mnist.py:58);
// Horovod 0.18.1 runs NCCL allreduce + divide + optimizer kernels after backward.
#pragma once
#include "comm.cuh"

// signal-pad slots used by the pipelined step (the bucket kernels use slots 0 .. n_buckets-1)
#define HVD_SLOT_RS_PUSH 24    // [tile-in-shard][src]  : src pushed this tile of my shard                (epoch value)
#define HVD_SLOT_TAIL 25       // epoch counters of the tail kernel's CTAs (no flags)
#define HVD_SLOT_READY 26      // [all-gather CTA][src] : src's CTA delivered its part of version v        (version value)

// GEMM epilogue "push": output tile j (128 rows x 128 columns, contiguous) goes to
//   recv[owner(j)] + rank * shard_elems + (j % tiles_per_owner) * 128 * 128      with owner(j) = j / tiles_per_owner
struct HvdRsPush {
  float* recv[HVD_MAX_RANKS];      // recv[r]: rank r's receive area as mapped in THIS process ([src][shard_elems])
  uint32_t* pad[HVD_MAX_RANKS];    // peer-mapped signal pads
  uint32_t* epoch;                 // this rank's private epoch counters (HVD_NUM_SLOTS x HVD_MAX_CTAS)
  int rank, nranks;
  int tiles_per_owner;
  long long shard_elems;
  int send_flags;                  // 1: each GEMM CTA raises its tile's flag itself; 0: hvd_rs_flags_launch does it
                                   //    for all tiles after the GEMM (the GEMM's CTAs then leave their SMs without
                                   //    sitting through a system fence)
};

// "The big weights are current" wait.  Every rank's all-gather CTA g stamps flag[g][src] on every peer with the
// version (number of updates applied) of the rows it delivered; a reader waits until all (src, g) flags have
// reached *version - the number of updates THIS rank has applied, which is the same on every rank because the
// tail kernels run in lock step.  Re-sending a version is harmless (idempotent), which keeps the host logic free
// of "is something pending" state shared between ranks.
struct HvdParamWait {
  const uint32_t* flags;           // this rank's pad + HVD_SLOT_READY block; nullptr: no wait
  const uint32_t* version;         // device word advanced by the tail kernel
  int* error;
  unsigned long long timeout_ns;
  int rank, nranks, ag_ctas;
  unsigned long long* trace;       // profiling ring (see pipeline_tail.cu), or nullptr: [12] wait began, [13] wait over
};

#define HVD_PARAM_WAIT_MAX_FLAGS 256     // nranks * ag_ctas <= 256: at most 8 flags per lane

__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Called by ONE WHOLE WARP (all 32 lanes).  Each lane keeps its (up to 8) flag loads in flight together - relaxed
// loads, re-issued until all have reached the version - and ONE acquire fence at the end orders the weight reads
// after them (an acquire load per flag would serialise 4-8 L2 round trips per lane).
__device__ __forceinline__ void hvd_param_wait_warp(const HvdParamWait& w) {
  if (w.flags == nullptr) return;
  const int lane = threadIdx.x & 31;
  const uint32_t v = *reinterpret_cast<const volatile uint32_t*>(w.version);
  const int total = w.nranks * w.ag_ctas;
  unsigned long long t0 = 0;
  unsigned spins = 0;
  unsigned long long* tr = (w.trace && lane == 0 && blockIdx.x + blockIdx.y + blockIdx.z == 0) ? w.trace + (size_t)(v & 63) * 16 : nullptr;
  if (tr) tr[12] = globaltimer_ns();
  for (;;) {
    bool all = true;
    int missing = 0;
#pragma unroll
    for (int j = 0; j < HVD_PARAM_WAIT_MAX_FLAGS / 32; j++) {
      const int k = lane + 32 * j;
      if (k < total) {
        const int r = k / w.ag_ctas, g = k - r * w.ag_ctas;
        if (r != w.rank && (int32_t)(ld_relaxed_sys_u32(w.flags + ((size_t)g * HVD_MAX_RANKS + r)) - v) < 0) { all = false; missing = r; }
      }
    }
    if (all) break;
    if (w.timeout_ns && ((++spins & 0xff) == 0)) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > w.timeout_ns) { atomicCAS(w.error, 0, (HVD_SLOT_READY << 8) | (missing + 1)); break; }
      if (*(volatile int*)w.error != 0) break;
    }
  }
  asm volatile("fence.acq_rel.sys;" ::: "memory");
  if (tr) tr[13] = globaltimer_ns();
  __syncwarp();
  // the weights were written by other GPUs' generic-proxy stores and are read next through TMA
  asm volatile("fence.proxy.async;" ::: "memory");
}

// ---------------------------------------------------------------------------------------------------------
// All-gather of the updated rows (step 3 above).  One "block" = a contiguous 1/nblk of this rank's rows, copied
// to every other replica with plain 16-byte loads and peer stores (or multimem.st), then one version flag per
// peer.  The same device function is (a) the stand-alone hvd_allgather_kernel used by flushes and (b) the role
// of the FIRST blocks of the conv1 forward kernel of the next step: blocks of one grid are dispatched in index
// order, so the copy starts with the step - a separate small kernel beside the conv forward was measured to
// wait until that kernel's CTAs had all retired, whatever its stream priority.
struct HvdAllgatherArgs {
  long long w0, w1;               // the big tensor's range in the slab
  long long param_off;            // slab offset of the parameters inside the arena buffer
  const uint32_t* version[HVD_MAX_RANKS];   // [own rank]: device word written by the tail kernel
  int use_mc;
  unsigned long long* trace[HVD_MAX_RANKS];  // same ring as the tail's: [10] block 0 started, [11] block 0 stamped its flags
};

// `peer`: a __shared__ float*[HVD_MAX_RANKS] of the calling kernel.  Contains __syncthreads(): call from every
// thread of the block.
__device__ __forceinline__ void hvd_allgather_block(const HvdCommCtx& c, const HvdAllgatherArgs& a, float** peer, int blk,
                                                    int nblk) {
  const int rank = hvd_rank(c), N = c.nranks, tid = threadIdx.x, nt = blockDim.x;
  const uint32_t v = *a.version[rank];
  unsigned long long* tr = (a.trace[rank] && blk == 0 && tid == 0) ? a.trace[rank] + (size_t)(v & 63) * 16 : nullptr;
  if (tr) tr[10] = globaltimer_ns();
  const long long shard = (a.w1 - a.w0) / N;
  const long long n4 = shard / 4;
  const long long per = (n4 + nblk - 1) / nblk;
  const long long i0 = (long long)blk * per, i1 = i0 + per < n4 ? i0 + per : n4;
  const long long off = a.param_off + a.w0 + (long long)rank * shard;
  const float* src = c.base[rank] + off;
  if (tid < HVD_MAX_RANKS) peer[tid] = c.base[tid];
  __syncthreads();
  for (long long i = i0 + tid; i < i1; i += 4 * nt) {
    float4 q[4];
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (i + u * nt < i1) q[u] = ld_f4(src + (i + u * nt) * 4);
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (i + u * nt >= i1) continue;
      const long long o = off + (i + u * nt) * 4;
      if (a.use_mc) {
        multimem_st_f4(c.mc_base + o, q[u]);          // also rewrites this rank's own copy with the same bits
      } else {
        for (int r = 0; r < N; r++)
          if (r != rank) st_sys_f4(peer[r] + o, q[u]);
      }
    }
  }
  // every thread's stores happen-before the barrier; the release of the signalling threads is cumulative
  __syncthreads();
  if (tid < N && tid != rank) st_release_sys(hvd_flag(c, tid, HVD_SLOT_READY, blk, rank), v);
  if (tr) tr[11] = globaltimer_ns();
}
