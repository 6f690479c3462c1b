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
