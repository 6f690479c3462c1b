// Tail kernel of the pipelined data-parallel step (comm/pipeline.cuh).
//
// By the time it runs, the big Dense-kernel gradient has already been reduce-scattered by the GEMM that
// produced it (tc_gemm EPI_RS_PUSH, hidden under the conv backward), so what is left on the critical path
// is small:
//
//   LL CTAs    (blockIdx.x <  ll_ctas)  the ~80 KB of small tensors (conv kernels, biases, Dense(10)).
//              One-shot allreduce with NO barrier: every rank pushes its values into every peer's receive
//              buffer as 16-byte packets {v0, flag, v1, flag} (flag = step epoch), then polls its own buffer
//              until each source's packet carries this step's flag, sums in fixed rank order (bit-identical
//              replicas), scales by 1/N and applies Adadelta / Adam / SGD to its own replica.  One NVLink
//              one-way latency instead of barrier + pull round trip + barrier.  Receive buffers are
//              double-buffered by step parity: a peer can be at most one step ahead (its step t+2 push needs
//              this rank's step t+1 push, which follows this rank's step t reads in stream order).
//   shard CTAs (blockIdx.x >= ll_ctas)  rank r owns rows [r*R/N, (r+1)*R/N) of the big kernel.  Its receive
//              area holds one slot per source rank, filled by the producing GEMMs.  Wait for the per-tile
//              flags (long since raised), sum the slots in fixed rank order, scale, update the owned rows of
//              this rank's parameters and of the optimizer state (state is sharded by owner).  Purely local
//              traffic, 1/N of the full update.
//
// hvd_allgather_kernel delivers the updated rows to the other replicas.  It is NOT part of the tail: a handful
// of 128-thread, 32-register CTAs enqueued at the start of the next step (or by a host-side flush) that fit on
// the SMs next to the conv forward CTAs, copy the owned rows into every peer's arena (multimem.st through the
// switch, or one peer store per rank) and stamp a version flag on every peer.  The 4+ MB that every GPU receives
// per step therefore crosses NVLink while the conv forward computes, and the first reader of the big weights
// (tc_gemm wait hook) checks the flags - nobody waits in an exit barrier and no kernel on the critical path sits
// through the acknowledgement of its peer stores (measured: 5 us at 2 ranks, 9 us at 8 when it was in the tail).
//
// Capability parity: the allreduce + 1/N + optimizer step of hvd.DistributedOptimizer (reference
// mnist_keras.py:84-87, tensorflow2_keras_mnist.py:55-58).
#include "pipeline.cuh"
#include "kernels/tc_common.cuh"

#include <cooperative_groups.h>
#include <string.h>

enum { HVD_KIND_NONE = 0, HVD_KIND_SGD = 1, HVD_KIND_ADADELTA = 2, HVD_KIND_ADAM = 3 };
enum { H_LR = 0, H_A = 1, H_B = 2, H_EPS = 3, H_STEP = 4, H_DECAY = 5, H_LR_T = 6 };

struct HvdTailArgs {
  // small tensors: two ranges of the slab (element offsets, multiples of 4), packed back to back
  long long a0, a1, b0, b1;
  // the big tensor [w0, w1): rows split evenly over the ranks ((w1 - w0) divisible by 4 * nranks)
  long long w0, w1;
  long long grad_off, param_off, slot0_off, slot1_off;   // slab offsets inside the arena buffer (c.base)
  float* recv[HVD_MAX_RANKS];     // recv[r]: rank r's receive area of the reduce-scatter, [src][shard]; only
                                  // [own rank] is read (the array form lets one emulated launch carry all ranks)
  float4* ll[HVD_MAX_RANKS];      // ll[r]: rank r's LL receive buffer as mapped here: [parity][src][packets]
  uint32_t* version[HVD_MAX_RANKS];        // [own rank]: device word = number of updates applied to the big tensor
                                  // (written here, stamped on the peers by the all-gather, awaited by readers)
  const float* hyper;             // device fp32[8]
  float scale;                    // 1/N
  int kind, parity, ll_ctas, use_mc;       // (use_mc: unused by the tail since the all-gather left it; kept for layout)
  // HVD_TIMELINE / profiling: ring of 64 records x 16 %globaltimer stamps (ns), indexed by epoch & 63, or null.
  //   [0] LL CTA 0 entered   [1] LL CTA 0 pushed its packets   [2] LL CTA 0 has every peer's packets   [3] LL CTA 0 done
  //   [4] first shard CTA entered  [5] its tiles were there  [7] it is done   ([10], [11]: hvd_allgather_kernel)
  unsigned long long* trace[HVD_MAX_RANKS];
};

template <int KIND>
__device__ __forceinline__ float tail_update_one(float g, float p, float& s0, float& s1, const float* __restrict__ h) {
  if (KIND == HVD_KIND_ADADELTA) {
    const float rho = h[H_A], eps = h[H_EPS];
    s0 = rho * s0 + (1.0f - rho) * g * g;
    const float u = g * sqrtf(s1 + eps) / sqrtf(s0 + eps);
    s1 = rho * s1 + (1.0f - rho) * u * u;
    return p - h[H_LR] * u;
  } else if (KIND == HVD_KIND_ADAM) {
    const float b1 = h[H_A], b2 = h[H_B];
    s0 = b1 * s0 + (1.0f - b1) * g;
    s1 = b2 * s1 + (1.0f - b2) * g * g;
    return p - h[H_LR_T] * s0 / (sqrtf(s1) + h[H_EPS]);
  } else if (KIND == HVD_KIND_SGD) {
    const float mom = h[H_A];
    s0 = mom * s0 - h[H_LR] * g;
    return (h[H_B] != 0.0f) ? p + mom * s0 - h[H_LR] * g : p + s0;
  }
  return g;
}

__device__ __forceinline__ float4 ld_volatile_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Bounded wait for `*flag >= epoch` (signed distance).  Returns false on timeout (error word raised).
__device__ __forceinline__ bool tail_wait_flag(const HvdCommCtx& c, const uint32_t* flag, uint32_t epoch, int slot, int peer) {
  unsigned long long t0 = 0;
  unsigned spins = 0;
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if (c.timeout_ns && ((++spins & 0x3ff) == 0)) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > c.timeout_ns) { atomicCAS((int*)c.error, 0, (slot << 8) | (peer + 1)); return false; }
      if (*(volatile int*)c.error != 0) return false;
    }
  }
  return true;
}

template <int KIND, int NR>
__global__ void __launch_bounds__(256, 2) hvd_tail_kernel(HvdCommCtx c, HvdTailArgs a) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int rank = hvd_rank(c);
  const int N = c.nranks;
  const int tid = threadIdx.x;
  const uint32_t ep = hvd_epoch_load(c, HVD_SLOT_TAIL) + 1;     // this step's epoch (same on every CTA and rank)
  float* mine = c.base[rank];
  float hloc[8];
#pragma unroll
  for (int k = 0; k < 8; k++) hloc[k] = a.hyper[k];
  __shared__ int s_fail;
  if (tid == 0) s_fail = 0;
  __syncthreads();
  unsigned long long* tr = a.trace[rank] ? a.trace[rank] + (size_t)(ep & 63) * 16 : nullptr;
  const bool tr_ll = tr && blockIdx.x == 0 && tid == 0;
  const bool tr_sh = tr && (int)blockIdx.x == a.ll_ctas && tid == 0;
  if (tr_ll) tr[0] = globaltimer_ns();
  if (tr_sh) tr[4] = globaltimer_ns();

  if ((int)blockIdx.x < a.ll_ctas) {
    // ------------------------------------------------------------------ small tensors: LL one-shot
    const long long na = a.a1 - a.a0, nsmall = na + (a.b1 - a.b0);
    const long long npk = nsmall / 2;                     // packets of two floats (ranges are multiples of 4)
    const uint32_t fl = ep;
    for (long long pk0 = (long long)blockIdx.x * blockDim.x; pk0 < npk; pk0 += (long long)a.ll_ctas * blockDim.x) {
      const long long pk = pk0 + tid;
      const bool on = pk < npk;
      long long el = 0;
      float2 g = make_float2(0.f, 0.f);
      if (on) {
        const long long i = 2 * pk;
        el = i < na ? a.a0 + i : a.b0 + (i - na);
        g = *reinterpret_cast<const float2*>(mine + a.grad_off + el);
        const float4 pkt = make_float4(g.x, __uint_as_float(fl), g.y, __uint_as_float(fl));
#pragma unroll
        for (int r = 0; r < NR; r++)
          if (r < N && r != rank) {
            float4* dst = a.ll[r] + ((long long)(a.parity * N + rank) * npk + pk);
            asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(pkt.x), "f"(pkt.y), "f"(pkt.z), "f"(pkt.w) : "memory");
          }
      }
      if (tr_ll && pk0 == 0) tr[1] = globaltimer_ns();
      // local parameter/state loads overlap the flight time of the peers' packets
      float2 p2 = make_float2(0.f, 0.f), s0 = p2, s1 = p2;
      if (on && KIND != HVD_KIND_NONE) {
        p2 = *reinterpret_cast<const float2*>(mine + a.param_off + el);
        s0 = *reinterpret_cast<const float2*>(mine + a.slot0_off + el);
        if (KIND != HVD_KIND_SGD) s1 = *reinterpret_cast<const float2*>(mine + a.slot1_off + el);
      }
      float2 acc = make_float2(0.f, 0.f);
      bool ok = true;
      if (on) {
#pragma unroll
        for (int r = 0; r < NR; r++) {                     // fixed order: identical bits on every rank
          if (r >= N) continue;
          float2 v = g;
          if (r != rank) {
            const float4* src = a.ll[rank] + ((long long)(a.parity * N + r) * npk + pk);
            float4 q = ld_volatile_f4(src);
            unsigned long long t0 = 0;
            unsigned spins = 0;
            while (__float_as_uint(q.y) != fl || __float_as_uint(q.w) != fl) {
              if (c.timeout_ns && ((++spins & 0x3ff) == 0)) {
                const unsigned long long now = globaltimer_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > c.timeout_ns) { atomicCAS((int*)c.error, 0, (HVD_SLOT_TAIL << 8) | (r + 1)); ok = false; break; }
                if (*(volatile int*)c.error != 0) { ok = false; break; }
              }
              q = ld_volatile_f4(src);
            }
            v = make_float2(q.x, q.z);
          }
          acc.x += v.x; acc.y += v.y;
        }
      }
      if (tr_ll && pk0 == 0) tr[2] = globaltimer_ns();
      if (on && ok) {
        acc.x *= a.scale; acc.y *= a.scale;
        if (KIND == HVD_KIND_NONE) {
          *reinterpret_cast<float2*>(mine + a.grad_off + el) = acc;
        } else {
          float2 res;
          res.x = tail_update_one<KIND>(acc.x, p2.x, s0.x, s1.x, hloc);
          res.y = tail_update_one<KIND>(acc.y, p2.y, s0.y, s1.y, hloc);
          *reinterpret_cast<float2*>(mine + a.slot0_off + el) = s0;
          if (KIND != HVD_KIND_SGD) *reinterpret_cast<float2*>(mine + a.slot1_off + el) = s1;
          *reinterpret_cast<float2*>(mine + a.param_off + el) = res;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ owned rows of the big tensor
    const int d = blockIdx.x - a.ll_ctas, D = gridDim.x - a.ll_ctas;
    const long long shard = (a.w1 - a.w0) / N;             // floats owned by one rank
    const long long n4 = shard / 4;
    const long long per = (n4 + D - 1) / D;
    const long long i0 = (long long)d * per, i1 = i0 + per < n4 ? i0 + per : n4;
    // the tiles of my shard that this CTA touches have been pushed by every source
    if (N > 1 && i0 < i1) {
      const int t0 = (int)((i0 * 4) / (128 * 128)), t1 = (int)((i1 * 4 - 1) / (128 * 128));
      for (int k = tid; k < (t1 - t0 + 1) * N; k += blockDim.x) {
        const int tl = t0 + k / N, r = k % N;
        if (r != rank && !tail_wait_flag(c, hvd_flag(c, rank, HVD_SLOT_RS_PUSH, tl, r), ep, HVD_SLOT_RS_PUSH, r)) s_fail = 1;
      }
    }
    __syncthreads();
    if (s_fail) return;
    if (tr_sh) tr[5] = globaltimer_ns();
    const long long el0 = a.w0 + (long long)rank * shard;   // first owned element in the slab
    const float* recv = a.recv[rank];
    constexpr int U = 2;
    for (long long base = i0; base < i1; base += (long long)blockDim.x * U) {
      const long long i = base + tid;
      float4 pv[U], s0v[U], s1v[U], acc[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const long long idx = i + (long long)u * blockDim.x;
        if (idx < i1) {
          const long long el = el0 + idx * 4;
          if (KIND != HVD_KIND_NONE) {
            pv[u] = ld_f4(mine + a.param_off + el);
            s0v[u] = ld_f4(mine + a.slot0_off + el);
            if (KIND != HVD_KIND_SGD) s1v[u] = ld_f4(mine + a.slot1_off + el);
          }
          float4 v[NR];
#pragma unroll
          for (int r = 0; r < NR; r++)
            if (r < N) v[r] = ld_sys_f4(recv + (long long)r * shard + idx * 4);   // written by peers: bypass L1
          acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int r = 0; r < NR; r++)                     // fixed order: the owner is the only one who sums
            if (r < N) { acc[u].x += v[r].x; acc[u].y += v[r].y; acc[u].z += v[r].z; acc[u].w += v[r].w; }
        }
      }
      float4 res[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const long long idx = i + (long long)u * blockDim.x;
        if (idx >= i1) continue;
        const long long el = el0 + idx * 4;
        float4 g = acc[u];
        g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale;
        if (KIND == HVD_KIND_NONE) { res[u] = g; continue; }
        res[u].x = tail_update_one<KIND>(g.x, pv[u].x, s0v[u].x, s1v[u].x, hloc);
        res[u].y = tail_update_one<KIND>(g.y, pv[u].y, s0v[u].y, s1v[u].y, hloc);
        res[u].z = tail_update_one<KIND>(g.z, pv[u].z, s0v[u].z, s1v[u].z, hloc);
        res[u].w = tail_update_one<KIND>(g.w, pv[u].w, s0v[u].w, s1v[u].w, hloc);
        *reinterpret_cast<float4*>(mine + a.slot0_off + el) = s0v[u];
        if (KIND != HVD_KIND_SGD) *reinterpret_cast<float4*>(mine + a.slot1_off + el) = s1v[u];
      }
      // the owner's own replica; the other replicas get these rows from hvd_allgather_kernel
      const long long dst_base = (KIND == HVD_KIND_NONE) ? a.grad_off : a.param_off;
#pragma unroll
      for (int u = 0; u < U; u++) {
        const long long idx = i + (long long)u * blockDim.x;
        if (idx < i1) *reinterpret_cast<float4*>(mine + dst_base + el0 + idx * 4) = res[u];
      }
    }
    if (tr_sh) tr[7] = globaltimer_ns();
    if (d == 0 && tid == 0) *a.version[rank] = ep;          // read by the all-gather / the readers' wait (stream order)
  }
  if (tr_ll) { tr[3] = globaltimer_ns(); tr[15] = ep; }      // [15]: which step this record belongs to
  hvd_epoch_store(c, HVD_SLOT_TAIL, ep);
}

// Wait (one warp) until every peer has delivered the current version of its rows into this arena.  Enqueued
// by the host before it reads parameters outside a training step (checkpoint, evaluate, get_weights) - a stream
// synchronize alone does not cover stores that OTHER GPUs are still sending.
__global__ void hvd_param_wait_kernel(HvdParamWait w) { hvd_param_wait_warp(w); }

// Deliver this rank's rows of the big tensor to every other replica and stamp the version on every peer.
// 128 threads, <= 32 registers: 4 K registers per CTA, which is what an SM has left beside a conv2 forward CTA.
__global__ void __launch_bounds__(128, 16) hvd_allgather_kernel(HvdCommCtx c, HvdAllgatherArgs a) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __shared__ float* peer[HVD_MAX_RANKS];       // shared memory, not registers: the kernel lives on 32 registers
  hvd_allgather_block(c, a, peer, blockIdx.x, gridDim.x);
}

// The same delivery with the TMA engine doing the copying: ONE WARP per CTA, one lane issuing bulk copies
// (cp.async.bulk global -> shared -> every peer's arena) through a ring of 16 KB shared-memory buffers.  No
// registers to speak of, no load latency to cover with threads: the version above needs thousands of threads to
// keep enough 16-byte loads in flight and took 9 us alone / 20-40 us squeezed in beside the conv forward CTAs
// with the 8 K threads that fit there; here a few dozen warps keep several hundred KB in flight.
constexpr int AG_CHUNK = 16384;
constexpr int AG_NBUF = 3;

__global__ void __launch_bounds__(32) hvd_allgather_tma_kernel(HvdCommCtx c, HvdAllgatherArgs a) {
  extern __shared__ __align__(128) uint8_t ag_smem[];
  __shared__ __align__(8) uint64_t full[AG_NBUF];
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int rank = hvd_rank(c), N = c.nranks, lane = threadIdx.x;
  const uint32_t v = *a.version[rank];
  unsigned long long* tr = (a.trace[rank] && blockIdx.x == 0 && lane == 0) ? a.trace[rank] + (size_t)(v & 63) * 16 : nullptr;
  if (tr) tr[10] = globaltimer_ns();
  const long long shard_bytes = (a.w1 - a.w0) / N * 4;
  long long per = (shard_bytes + gridDim.x - 1) / gridDim.x;
  per = (per + 127) / 128 * 128;
  const long long b0 = (long long)blockIdx.x * per;
  const long long b1 = b0 + per < shard_bytes ? b0 + per : shard_bytes;
  const long long off = (a.param_off + a.w0) * 4 + (long long)rank * shard_bytes;      // byte offset in the arena buffer
  if (lane == 0) {
    for (int i = 0; i < AG_NBUF; i++) tc::mbar_init(&full[i], 1);
    tc::fence_barrier_init();
    const int nchunk = b1 > b0 ? (int)((b1 - b0 + AG_CHUNK - 1) / AG_CHUNK) : 0;
    const char* src = reinterpret_cast<const char*>(c.base[rank]) + off;
    auto load = [&](int k) {
      const long long o = b0 + (long long)k * AG_CHUNK;
      const uint32_t bytes = (uint32_t)((b1 - o) < AG_CHUNK ? (b1 - o) : AG_CHUNK);
      uint64_t* bar = &full[k % AG_NBUF];
      tc::mbar_expect_tx(bar, bytes);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(tc::smem_u32(ag_smem + (k % AG_NBUF) * AG_CHUNK)), "l"(src + o), "r"(bytes), "r"(tc::smem_u32(bar)) : "memory");
    };
    for (int k = 0; k < AG_NBUF - 1 && k < nchunk; k++) load(k);
    for (int k = 0; k < nchunk; k++) {
      tc::mbar_wait(&full[k % AG_NBUF], (k / AG_NBUF) & 1);
      const long long o = b0 + (long long)k * AG_CHUNK;
      const uint32_t bytes = (uint32_t)((b1 - o) < AG_CHUNK ? (b1 - o) : AG_CHUNK);
      for (int r = 0; r < N; r++) {
        if (r == rank) continue;
        char* dst = reinterpret_cast<char*>(c.base[r]) + off + o;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     ::"l"(dst), "r"(tc::smem_u32(ag_smem + (k % AG_NBUF) * AG_CHUNK)), "r"(bytes) : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (k + AG_NBUF - 1 < nchunk) {
        // the buffer chunk k-1 was sent from is the one chunk k+NBUF-1 lands in: its stores must have read it
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        load(k + AG_NBUF - 1);
      }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // every peer store of this CTA is complete ...
    asm volatile("fence.proxy.async;" ::: "memory");               // ... and ordered before the generic-proxy flag stores
  }
  __syncwarp();
  if (lane < N && lane != rank) st_release_sys(hvd_flag(c, lane, HVD_SLOT_READY, blockIdx.x, rank), v);
  if (tr) tr[11] = globaltimer_ns();
}

// Raise the "tile has landed" flags of a finished EPI_RS_PUSH GEMM (HvdRsPush.send_flags == 0): launched right
// behind the GEMM on the same stream, so the GEMM's stores are complete when this grid starts and the
// release below (cumulative over everything that happened before) publishes them.  One thread per tile.
__global__ void hvd_rs_flags_kernel(HvdRsPush rs, int ntiles) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int tile = blockIdx.x * blockDim.x + threadIdx.x;
  if (tile >= ntiles) return;
  uint32_t* ep = rs.epoch + HVD_SLOT_RS_PUSH * HVD_MAX_CTAS + tile;
  const uint32_t e = *ep + 1;
  const int owner = tile / rs.tiles_per_owner, tl = tile - owner * rs.tiles_per_owner;
  if (owner != rs.rank)
    st_release_sys(rs.pad[owner] + ((size_t)(HVD_SLOT_RS_PUSH * HVD_MAX_CTAS + tl) * HVD_MAX_RANKS + rs.rank), e);
  *ep = e;
}

extern "C" int hvd_rs_flags_launch(const HvdRsPush* rs, void* stream) {
  if (!rs || rs->nranks < 1 || rs->nranks > HVD_MAX_RANKS) return -1;
  if (rs->nranks == 1) return 0;
  const int ntiles = rs->tiles_per_owner * rs->nranks;
  hvd_rs_flags_kernel<<<(ntiles + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*rs, ntiles);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------ launchers
template <int K, int NR>
static cudaError_t tail_launch_kn(const HvdCommCtx& c, const HvdTailArgs& a, int ctas, cudaStream_t st, int emu) {
  if (emu > 0) {
    void* args[] = {(void*)&c, (void*)&a};
    return cudaLaunchCooperativeKernel((void*)hvd_tail_kernel<K, NR>, dim3(ctas, emu), dim3(256), args, 0, st);
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(256); cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, hvd_tail_kernel<K, NR>, c, a);
}

template <int K>
static cudaError_t tail_launch_k(const HvdCommCtx& c, const HvdTailArgs& a, int ctas, cudaStream_t st, int emu) {
  if (c.nranks <= 1) return tail_launch_kn<K, 1>(c, a, ctas, st, emu);
  if (c.nranks <= 2) return tail_launch_kn<K, 2>(c, a, ctas, st, emu);
  if (c.nranks <= 4) return tail_launch_kn<K, 4>(c, a, ctas, st, emu);
  return tail_launch_kn<K, 8>(c, a, ctas, st, emu);
}

// emulate_ranks > 0: ONE cooperative launch, blockIdx.y is the rank (tests on a single GPU).
extern "C" int hvd_tail_launch(const HvdCommCtx* c, const HvdTailArgs* a, int ctas, void* stream, int emulate_ranks) {
  if (!c || !a || c->nranks < 1 || c->nranks > HVD_MAX_RANKS) return -1;
  if (ctas < 2 || ctas > HVD_MAX_CTAS || a->ll_ctas < 1 || a->ll_ctas >= ctas) return -2;
  if ((a->a0 | a->a1 | a->b0 | a->b1 | a->w0 | a->w1 | a->grad_off | a->param_off | a->slot0_off | a->slot1_off) & 3) return -3;
  if ((a->w1 - a->w0) % (4 * c->nranks) != 0 || ((a->w1 - a->w0) / c->nranks) % (128 * 128) != 0) return -4;
  cudaStream_t st = (cudaStream_t)stream;
  const int emu = emulate_ranks;
  cudaError_t e;
  switch (a->kind) {
    case HVD_KIND_NONE: e = tail_launch_k<HVD_KIND_NONE>(*c, *a, ctas, st, emu); break;
    case HVD_KIND_SGD: e = tail_launch_k<HVD_KIND_SGD>(*c, *a, ctas, st, emu); break;
    case HVD_KIND_ADADELTA: e = tail_launch_k<HVD_KIND_ADADELTA>(*c, *a, ctas, st, emu); break;
    case HVD_KIND_ADAM: e = tail_launch_k<HVD_KIND_ADAM>(*c, *a, ctas, st, emu); break;
    default: return -7;
  }
  return (int)e;
}

// One %globaltimer reading (ns) into *out - lets the host line device stamps up with its own clock.
__global__ void hvd_globaltimer_kernel(unsigned long long* out) { *out = globaltimer_ns(); }
extern "C" int hvd_globaltimer_probe(unsigned long long* out, void* stream) {
  hvd_globaltimer_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(out);
  return (int)cudaGetLastError();
}

extern "C" int hvd_param_wait_launch(const HvdParamWait* w, void* stream) {
  if (!w) return -1;
  hvd_param_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*w);
  return (int)cudaGetLastError();
}

// emulate_ranks > 0: grid.y = rank (a plain launch: the kernel waits for nobody).
// use_tma: the bulk-copy kernel (peer stores only; ignores a->use_mc).
extern "C" int hvd_allgather_launch(const HvdCommCtx* c, const HvdAllgatherArgs* a, int ctas, void* stream,
                                    int emulate_ranks, int use_tma) {
  if (!c || !a || c->nranks < 2 || c->nranks > HVD_MAX_RANKS || ctas < 1 || ctas > HVD_MAX_CTAS) return -1;
  if ((a->w0 | a->w1 | a->param_off) & 3 || (a->w1 - a->w0) % (4 * c->nranks) != 0) return -2;
  if (a->use_mc && c->mc_base == nullptr) return -3;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(ctas, emulate_ranks > 0 ? emulate_ranks : 1); cfg.blockDim = dim3(128); cfg.stream = (cudaStream_t)stream;
  // Highest priority: CTAs are dispatched grid by grid in submission order within a priority level, so without it
  // these CTAs queue behind the conv forward kernels' not-yet-resident CTAs instead of slipping into the free
  // thread slots beside them (measured: the copy started 7 us late and took 30 us).  Kept by graph capture.
  int lo = 0, hi = 0;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributePriority;
  at[0].val.priority = hi;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (use_tma) {
    constexpr int smem = AG_NBUF * AG_CHUNK;
    static tc::PerDeviceOnce attr;
    if (attr.need()) {
      if (cudaFuncSetAttribute(hvd_allgather_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -10;
      attr.mark();
    }
    cfg.blockDim = dim3(32);
    cfg.dynamicSmemBytes = smem;
    return (int)cudaLaunchKernelEx(&cfg, hvd_allgather_tma_kernel, *c, *a);
  }
  return (int)cudaLaunchKernelEx(&cfg, hvd_allgather_kernel, *c, *a);
}

// sizeof of the argument structs, in the order HvdCommCtx, HvdRsPush, HvdParamWait, HvdTailArgs, HvdAllgatherArgs:
// the ctypes mirrors (parallel/peer.py, parallel/pipeline.py) are checked against these in the CPU tests.
extern "C" void hvd_pipeline_struct_sizes(int* out5) {
  out5[0] = (int)sizeof(HvdCommCtx);
  out5[1] = (int)sizeof(HvdRsPush);
  out5[2] = (int)sizeof(HvdParamWait);
  out5[3] = (int)sizeof(HvdTailArgs);
  out5[4] = (int)sizeof(HvdAllgatherArgs);
}
