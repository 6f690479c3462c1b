// Descriptors of the pipelined data-parallel step: the gradient of the big
// Dense kernel never sits in a local gradient buffer waiting for a collective - the GEMM that produces
// it pushes every output tile straight into the receive area of the rank that owns those rows
// (GEMM -> reduce-scatter in one kernel, over peer memory), the step's tail kernel reduces + updates the
// owned rows and stores the new weights into every replica, and the first kernel of the NEXT step that
// reads those weights waits on a counter instead of the tail kernel waiting in an exit barrier.
//
// Capability parity: hvd.DistributedOptimizer (reference mnist_keras.py:87, tensorflow2_keras_mnist.py:58);
// Horovod 0.18.1 runs NCCL allreduce + divide + optimizer kernels after backward.
#pragma once
#include "comm.cuh"

// signal-pad slots used by the pipelined step (the bucket kernels use slots 0 .. n_buckets-1)
#define HVD_SLOT_RS_PUSH 24    // [tile-in-shard][src]  : src pushed this tile of my shard           (epoch value)
#define HVD_SLOT_TAIL 25       // [0][src]              : src entered its tail kernel of this step   (epoch value)
#define HVD_SLOT_READY 26      // [0][0]                : counter, +1 per tail CTA of every rank whose stores landed

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

// "Parameters are ready" wait, executed by the first consumer of weights that peers store into this
// rank's arena: spin until *counter >= *target (both device words; the target is advanced by the tail kernel).
struct HvdParamWait {
  const uint32_t* counter;
  const uint32_t* target;
  int* error;
  unsigned long long timeout_ns;
};

__device__ __forceinline__ void hvd_param_wait(const HvdParamWait& w) {
  if (w.counter == nullptr) return;
  const uint32_t tgt = *reinterpret_cast<const volatile uint32_t*>(w.target);
  unsigned long long t0 = 0;
  unsigned spins = 0;
  while ((int32_t)(ld_acquire_sys(w.counter) - tgt) < 0) {
    if (w.timeout_ns && ((++spins & 0x3ff) == 0)) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > w.timeout_ns) { atomicCAS(w.error, 0, (HVD_SLOT_READY << 8) | 0xff); break; }
    }
  }
  // the weights were written by other GPUs' generic-proxy stores and are read next through TMA
  asm volatile("fence.proxy.async;" ::: "memory");
}
