// Cross-GPU primitives for the peer-memory collectives (sm_100a).
//
// Every rank maps every peer's symmetric buffer (cuMem VMM or CUDA IPC) and a small
// signal pad.  Kernels synchronise with monotonically increasing epoch counters written
// with st.release.sys into the peers' pads and polled with ld.acquire.sys on the local pad
// (no 0/1 flags -> no ABA under CUDA-graph replay).  Spins are
// bounded: on timeout the kernel raises an error word in host-visible memory and ABORTS - it
// returns without reducing, updating, storing to peers or advancing its epoch, so parameters and
// optimizer state stay as they were when the host reads the error word (stall inspector
// equivalent).  The bound is the shutdown time, not the 60 s warning time.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define HVD_MAX_RANKS 8
#define HVD_MAX_CTAS 320
#define HVD_NUM_SLOTS 32

struct HvdCommCtx {
  int rank;                              // >= 0: this process' rank.  -1: emulation, rank = blockIdx.y
  int nranks;
  float* base[HVD_MAX_RANKS];            // peer-mapped symmetric buffer of each rank
  float* mc_base;                        // multicast (NVLS) mapping of the same buffer, or nullptr
  uint32_t* pad[HVD_MAX_RANKS];          // peer-mapped signal pad of each rank
  uint32_t* epoch[HVD_MAX_RANKS];        // per-rank private epoch counters (only [rank] used in production)
  int* error;                            // host-visible error word (0 = ok)
  unsigned long long timeout_ns;         // 0 = spin forever
};

__device__ __forceinline__ int hvd_rank(const HvdCommCtx& c) { return c.rank >= 0 ? c.rank : (int)blockIdx.y; }

__device__ __forceinline__ uint32_t* hvd_flag(const HvdCommCtx& c, int owner, int slot, int cta, int src) {
  return c.pad[owner] + ((size_t)(slot * HVD_MAX_CTAS + cta) * HVD_MAX_RANKS + src);
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Coherent 16-byte peer load / store (bypass the non-coherent L1 for remote data).
__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
  float4 v;
  // no "memory" clobber: ordering against the barriers comes from their own fences, and leaving it
  // out lets the compiler keep many of these in flight
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st_sys_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// NVLS: the switch returns the sum over all ranks' copies / broadcasts the store to all.
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc));
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Per-CTA cross-rank barrier.  CTA b of every rank signals CTA b of every peer and waits
// for all of them: no grid-wide sync, and the only co-residency requirement is that CTA b
// of each rank eventually runs (grid <= resident capacity, enforced by the launcher).
//   release=true : make this CTA's earlier global/peer stores visible first.
// Returns false (for the whole CTA) when a peer did not arrive within timeout_ns, or when another CTA /
// an earlier kernel already raised the error word: the caller must return without touching anything.
__device__ __forceinline__ bool hvd_barrier(const HvdCommCtx& c, int slot, uint32_t epoch, bool release) {
  const int rank = hvd_rank(c);
  const int cta = blockIdx.x;
  if (c.nranks == 1) { __syncthreads(); return true; }   // single GPU: nothing to wait for
  if (release) __threadfence_system();
  __syncthreads();
  int failed = 0;
  if ((int)threadIdx.x < c.nranks) {
    const int peer = threadIdx.x;
    st_release_sys(hvd_flag(c, peer, slot, cta, rank), epoch);
    const uint32_t* mine = hvd_flag(c, rank, slot, cta, peer);
    unsigned long long t0 = 0;
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (c.timeout_ns && ((++spins & 0x3ff) == 0)) {
        unsigned long long now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > c.timeout_ns) {
          // error word: 1-based missing peer in the low byte, slot above it
          atomicCAS((int*)c.error, 0, (slot << 8) | (peer + 1));
          failed = 1;
          break;
        }
        if (*(volatile int*)c.error != 0) { failed = 1; break; }     // somebody else already gave up
      }
    }
  }
  return __syncthreads_or(failed) == 0;
}

__device__ __forceinline__ uint32_t hvd_epoch_load(const HvdCommCtx& c, int slot) {
  return c.epoch[hvd_rank(c)][slot * HVD_MAX_CTAS + blockIdx.x];
}
__device__ __forceinline__ void hvd_epoch_store(const HvdCommCtx& c, int slot, uint32_t v) {
  __syncthreads();
  if (threadIdx.x == 0) c.epoch[hvd_rank(c)][slot * HVD_MAX_CTAS + blockIdx.x] = v;
}
