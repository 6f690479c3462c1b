// Fused peer-memory collectives for sm_100a: allreduce + 1/N scale + optimizer update in ONE
// kernel, reading and writing peer GPUs' arenas over NVLink 5 / NVSwitch from inside the kernel.
//
// Capability parity: this is the data path behind hvd.DistributedOptimizer (reference
// mnist_keras.py:87, tensorflow2_keras_mnist.py:58), BroadcastGlobalVariablesCallback
// (mnist_keras.py:97) and MetricAverageCallback (tensorflow2_keras_mnist.py:77).  Horovod 0.18.1
// does memcpy-in -> ncclAllReduce(ring) -> memcpy-out -> divide -> ~10 optimizer kernels per
// weight; here a gradient bucket is a contiguous arena range and each step is one launch.
//
// Variants (picked per message size from the measured table, parallel/variants.py):
//   ONESHOT  every rank reads all N ranks' bucket, sums in fixed rank order (bit-identical on
//            all ranks), updates its own replica.  (N-1)*S inbound; one barrier; latency-optimal.
//   TWOSHOT  rank r owns shard r: reduce-scatter by peer loads, update the shard (optimizer state
//            is sharded for these buckets), all-gather the UPDATED PARAMS by peer stores.
//            2*S*(N-1)/N per GPU.
//   NVLS     as TWOSHOT but the switch reduces (multimem.ld_reduce) and broadcasts (multimem.st):
//            ~S/N in and S/N out per GPU.
#include "comm.cuh"

#include <cooperative_groups.h>
#include <stdio.h>
#include <string.h>

enum { HVD_ONESHOT = 0, HVD_TWOSHOT = 1, HVD_NVLS = 2 };
enum { HVD_KIND_NONE = 0, HVD_KIND_SGD = 1, HVD_KIND_ADADELTA = 2, HVD_KIND_ADAM = 3 };
// hyper vector layout (keras/optimizers.py)
enum { H_LR = 0, H_A = 1, H_B = 2, H_EPS = 3, H_STEP = 4, H_DECAY = 5, H_LR_T = 6 };

struct HvdFusedArgs {
  long long start, end;                                   // element range in the slab (multiples of 4)
  long long grad_off, param_off, slot0_off, slot1_off;    // slab offsets inside the symmetric buffer
  long long out_off;                                      // KIND_NONE: where the averaged result goes
  const float* hyper;                                     // device fp32[8]
  float scale;                                            // 1/N (average) or 1 (sum)
  int kind, variant, slot, exit_barrier;
};

template <int KIND>
__device__ __forceinline__ float update_one(float g, float p, float& s0, float& s1, const float* __restrict__ h) {
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

template <int KIND>
__device__ __forceinline__ float4 update_f4(float4 g, float4 p, float4& s0, float4& s1, const float* __restrict__ h) {
  p.x = update_one<KIND>(g.x, p.x, s0.x, s1.x, h);
  p.y = update_one<KIND>(g.y, p.y, s0.y, s1.y, h);
  p.z = update_one<KIND>(g.z, p.z, s0.z, s1.z, h);
  p.w = update_one<KIND>(g.w, p.w, s0.w, s1.w, h);
  return p;
}

// NR = compile-time bound on the rank count (1, 2, 4, 8).  Fewer peers leave registers for a deeper
// unroll: the kernel is a latency-bound stream, so every thread keeps as many 16-byte loads in flight
// as its register budget allows (NVLS needs one load per element whatever the rank count).
template <int VARIANT, int NR>
struct Unroll { static constexpr int value = (VARIANT == HVD_NVLS || NR <= 2) ? 4 : 2; };

// <= 128 registers per thread (two CTAs per SM) on purpose: in the training step this kernel overlaps
// the conv backward kernels, whose CTAs pin every SM's shared memory; it only gets on an SM if its
// register footprint fits beside them.
template <int VARIANT, int KIND, int NR>
__global__ void __launch_bounds__(256, 2) hvd_fused_kernel(HvdCommCtx c, HvdFusedArgs a) {
  constexpr int kUnroll = Unroll<VARIANT, NR>::value;
  // programmatic dependent launch: this grid may have been scheduled while the previous kernel in the
  // stream was still draining; nothing above touches memory
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int rank = hvd_rank(c);
  const int N = c.nranks;
  const uint32_t ep = hvd_epoch_load(c, a.slot);
  // entry barrier: every rank's backward has finished writing this bucket
  if (!hvd_barrier(c, a.slot, ep + 1, false)) return;     // timed out: abort with params/state/epoch untouched

  long long s = a.start, e = a.end;
  if (VARIANT != HVD_ONESHOT) {
    long long per = (((a.end - a.start) / 4 + N - 1) / N) * 4;
    s = a.start + (long long)rank * per;
    e = s + per < a.end ? s + per : a.end;
    if (s > e) s = e;
  }
  const long long n4 = (e - s) / 4;
  const long long per_cta = (n4 + gridDim.x - 1) / gridDim.x;
  const long long i0 = (long long)blockIdx.x * per_cta;
  const long long i1 = i0 + per_cta < n4 ? i0 + per_cta : n4;
  float* mine = c.base[rank];
  float hloc[8];
  if (KIND != HVD_KIND_NONE) {
#pragma unroll
    for (int k = 0; k < 8; k++) hloc[k] = a.hyper[k];
  }

  for (long long i = i0 + threadIdx.x; i < i1; i += (long long)blockDim.x * kUnroll) {
    float4 acc[kUnroll];
    // local parameter/state loads are issued FIRST so they overlap the (much slower) peer loads
    float4 pv[kUnroll], s0v[kUnroll], s1v[kUnroll];
    if (KIND != HVD_KIND_NONE) {
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const long long idx = i + (long long)u * blockDim.x;
        if (idx < i1) {
          const long long el = s + idx * 4;
          pv[u] = ld_f4(mine + a.param_off + el);
          s0v[u] = ld_f4(mine + a.slot0_off + el);
          if (KIND != HVD_KIND_SGD) s1v[u] = ld_f4(mine + a.slot1_off + el);
        }
      }
    }
    if (VARIANT == HVD_NVLS) {
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const long long idx = i + (long long)u * blockDim.x;
        if (idx < i1) acc[u] = multimem_ld_reduce_f4(c.mc_base + a.grad_off + s + idx * 4);
      }
    } else {
      float4 v[kUnroll][NR];
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        const long long idx = i + (long long)u * blockDim.x;
#pragma unroll
        for (int r = 0; r < NR; r++)
          if (r < N && idx < i1)          // own gradients: plain load (written by earlier kernels on this GPU)
            v[u][r] = (r == rank) ? ld_f4(c.base[r] + a.grad_off + s + idx * 4)
                                  : ld_sys_f4(c.base[r] + a.grad_off + s + idx * 4);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; u++) {
        acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < NR; r++)                 // fixed order: identical bits on every rank
          if (r < N) { acc[u].x += v[u][r].x; acc[u].y += v[u][r].y; acc[u].z += v[u][r].z; acc[u].w += v[u][r].w; }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; u++) {
      const long long idx = i + (long long)u * blockDim.x;
      if (idx >= i1) continue;
      const long long el = s + idx * 4;
      float4 g = acc[u];
      g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale;
      float4 res;
      long long dst_off;
      if (KIND == HVD_KIND_NONE) { res = g; dst_off = a.out_off + el; }
      else {
        res = update_f4<KIND>(g, pv[u], s0v[u], s1v[u], hloc);
        *reinterpret_cast<float4*>(mine + a.slot0_off + el) = s0v[u];
        if (KIND != HVD_KIND_SGD) *reinterpret_cast<float4*>(mine + a.slot1_off + el) = s1v[u];
        dst_off = a.param_off + el;
      }
      if (VARIANT == HVD_ONESHOT) {
        *reinterpret_cast<float4*>(mine + dst_off) = res;
      } else if (VARIANT == HVD_TWOSHOT) {
#pragma unroll
        for (int r = 0; r < NR; r++)
          if (r < N) {
            if (r == rank) *reinterpret_cast<float4*>(mine + dst_off) = res;
            else st_sys_f4(c.base[r] + dst_off, res);
          }
      } else {
        multimem_st_f4(c.mc_base + dst_off, res);
      }
    }
  }

  uint32_t next = ep + 1;
  if (VARIANT != HVD_ONESHOT || a.exit_barrier) {
    // exit barrier: (two-shot/NVLS) every owner's stores have landed in every arena;
    // (one-shot) every rank has finished reading my gradients.
    if (!hvd_barrier(c, a.slot, ep + 2, VARIANT != HVD_ONESHOT)) return;
    next = ep + 2;
  }
  hvd_epoch_store(c, a.slot, next);
}

// ------------------------------------------------------------------------------ LL one-shot (small messages)
// Latency-optimal allreduce for messages up to a few hundred KB (the packed metric vector of
// MetricAverageCallback, reference tensorflow2_keras_mnist.py:77; small tensors of any caller): NO barrier at all.
// Every rank pushes its values into every peer's receive buffer as 16-byte packets {v0, flag, v1, flag} whose
// flag is the launch epoch, then polls its own buffer until each source's packet carries this epoch, sums in
// fixed rank order (bit-identical results), scales and writes.  Cost: one NVLink one-way latency instead of
// barrier + pull round trip (+ barrier).  Receive buffers are double-buffered by epoch parity: a peer can be at
// most one collective ahead (its launch k+2 needs this rank's launch k+1 push, which follows this rank's launch
// k reads in stream order).  Safe in place.  Relies, like NCCL's LL protocol, on 8-byte stores arriving whole.
#define HVD_SLOT_LL 27
struct HvdLLArgs {
  long long in_off, out_off, n;          // element offsets in the symmetric buffer; n even
  float scale;
  float4* ll[HVD_MAX_RANKS];             // ll[r]: rank r's receive buffer as mapped here: [parity][src][cap packets]
  long long cap;                         // packets per (parity, src)
};

template <int NR>
__global__ void __launch_bounds__(256) hvd_ll_allreduce_kernel(HvdCommCtx c, HvdLLArgs a) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int rank = hvd_rank(c), N = c.nranks;
  const uint32_t ep = hvd_epoch_load(c, HVD_SLOT_LL) + 1;
  const int par = ep & 1;
  const long long pk = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long npk = a.n / 2;
  float* mine = c.base[rank];
  if (pk < npk) {
    const float2 g = *reinterpret_cast<const float2*>(mine + a.in_off + 2 * pk);
    const float fl = __uint_as_float(ep);
#pragma unroll
    for (int r = 0; r < NR; r++)
      if (r < N && r != rank) {
        float4* dst = a.ll[r] + ((long long)(par * N + rank) * a.cap + pk);
        asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(g.x), "f"(fl), "f"(g.y), "f"(fl) : "memory");
      }
    float2 acc = make_float2(0.f, 0.f);
    bool ok = true;
#pragma unroll
    for (int r = 0; r < NR; r++) {                 // fixed order: identical bits on every rank
      if (r >= N) continue;
      float2 v = g;
      if (r != rank) {
        const float4* src = a.ll[rank] + ((long long)(par * N + r) * a.cap + pk);
        float4 q;
        unsigned long long t0 = 0;
        unsigned spins = 0;
        for (;;) {
          asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w) : "l"(src) : "memory");
          if (__float_as_uint(q.y) == ep && __float_as_uint(q.w) == ep) break;
          if (c.timeout_ns && ((++spins & 0x3ff) == 0)) {
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > c.timeout_ns) { atomicCAS((int*)c.error, 0, (HVD_SLOT_LL << 8) | (r + 1)); ok = false; break; }
            if (*(volatile int*)c.error != 0) { ok = false; break; }
          }
        }
        v = make_float2(q.x, q.z);
      }
      acc.x += v.x; acc.y += v.y;
    }
    if (ok) *reinterpret_cast<float2*>(mine + a.out_off + 2 * pk) = make_float2(acc.x * a.scale, acc.y * a.scale);
  }
  hvd_epoch_store(c, HVD_SLOT_LL, ep);
}

extern "C" int hvd_ll_allreduce_launch(const HvdCommCtx* c, const HvdLLArgs* a, void* stream, int emulate_ranks) {
  if (!c || !a || c->nranks < 1 || c->nranks > HVD_MAX_RANKS) return -1;
  if (a->n < 2 || (a->n & 1) || ((a->in_off | a->out_off) & 1) || a->n / 2 > a->cap) return -2;
  const int ctas = (int)((a->n / 2 + 255) / 256);
  if (ctas > HVD_MAX_CTAS) return -3;
  cudaStream_t st = (cudaStream_t)stream;
  void* fn;
  if (c->nranks <= 2) fn = (void*)hvd_ll_allreduce_kernel<2>;
  else if (c->nranks <= 4) fn = (void*)hvd_ll_allreduce_kernel<4>;
  else fn = (void*)hvd_ll_allreduce_kernel<8>;
  void* args[] = {(void*)c, (void*)a};
  if (emulate_ranks > 0)
    return (int)cudaLaunchCooperativeKernel(fn, dim3(ctas, emulate_ranks), dim3(256), args, 0, st);
  return (int)cudaLaunchKernel(fn, dim3(ctas), dim3(256), args, 0, st);
}

// Broadcast [off, off+n) from root's buffer into every rank's buffer (pull over NVLink).
__global__ void __launch_bounds__(512) hvd_bcast_kernel(HvdCommCtx c, int slot, int root, long long off, long long n) {
  const int rank = hvd_rank(c);
  const uint32_t ep = hvd_epoch_load(c, slot);
  if (!hvd_barrier(c, slot, ep + 1, false)) return;
  if (rank != root) {
    const long long n4 = n / 4;
    const float* src = c.base[root] + off;
    float* dst = c.base[rank] + off;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
      *reinterpret_cast<float4*>(dst + i * 4) = ld_sys_f4(src + i * 4);
  }
  if (!hvd_barrier(c, slot, ep + 2, false)) return;     // root may modify its data again
  hvd_epoch_store(c, slot, ep + 2);
}

// ------------------------------------------------------------------------------ launchers
template <int V, int K, int NR>
static cudaError_t launch_vkn(const HvdCommCtx& c, const HvdFusedArgs& a, int ctas, int threads, cudaStream_t st, int emu) {
  if (emu > 0) {
    void* args[] = {(void*)&c, (void*)&a};
    return cudaLaunchCooperativeKernel((void*)hvd_fused_kernel<V, K, NR>, dim3(ctas, emu), dim3(threads), args, 0, st);
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(threads); cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, hvd_fused_kernel<V, K, NR>, c, a);
}

template <int V, int K>
static cudaError_t launch_vk(const HvdCommCtx& c, const HvdFusedArgs& a, int ctas, int threads, cudaStream_t st, int emu) {
  if (c.nranks <= 1) return launch_vkn<V, K, 1>(c, a, ctas, threads, st, emu);
  if (c.nranks <= 2) return launch_vkn<V, K, 2>(c, a, ctas, threads, st, emu);
  if (c.nranks <= 4) return launch_vkn<V, K, 4>(c, a, ctas, threads, st, emu);
  return launch_vkn<V, K, 8>(c, a, ctas, threads, st, emu);
}

template <int V>
static cudaError_t launch_v(const HvdCommCtx& c, const HvdFusedArgs& a, int ctas, int threads, cudaStream_t st, int emu) {
  switch (a.kind) {
    case HVD_KIND_NONE: return launch_vk<V, HVD_KIND_NONE>(c, a, ctas, threads, st, emu);
    case HVD_KIND_SGD: return launch_vk<V, HVD_KIND_SGD>(c, a, ctas, threads, st, emu);
    case HVD_KIND_ADADELTA: return launch_vk<V, HVD_KIND_ADADELTA>(c, a, ctas, threads, st, emu);
    case HVD_KIND_ADAM: return launch_vk<V, HVD_KIND_ADAM>(c, a, ctas, threads, st, emu);
  }
  return cudaErrorInvalidValue;
}

static int check_args(const HvdCommCtx* c, const HvdFusedArgs* a, int ctas, int threads) {
  if (!c || !a || c->nranks < 1 || c->nranks > HVD_MAX_RANKS) return 1;
  if (ctas < 1 || ctas > HVD_MAX_CTAS || threads < 32 || threads > 256 || threads < c->nranks) return 2;
  if (a->slot < 0 || a->slot >= HVD_NUM_SLOTS) return 3;
  if ((a->start | a->end | a->grad_off | a->param_off | a->slot0_off | a->slot1_off | a->out_off) & 3) return 4;
  if (a->variant == HVD_NVLS && c->mc_base == nullptr) return 5;
  return 0;
}

extern "C" int hvd_fused_allreduce_launch(const HvdCommCtx* c, const HvdFusedArgs* a, int ctas, int threads,
                                          void* stream, int emulate_ranks) {
  int bad = check_args(c, a, ctas, threads);
  if (bad) return -bad;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  switch (a->variant) {
    case HVD_ONESHOT: e = launch_v<HVD_ONESHOT>(*c, *a, ctas, threads, st, emulate_ranks); break;
    case HVD_TWOSHOT: e = launch_v<HVD_TWOSHOT>(*c, *a, ctas, threads, st, emulate_ranks); break;
    case HVD_NVLS: e = launch_v<HVD_NVLS>(*c, *a, ctas, threads, st, emulate_ranks); break;
    default: return -6;
  }
  return (int)e;
}

extern "C" int hvd_bcast_launch(const HvdCommCtx* c, int slot, int root, long long off, long long n, int ctas,
                                int threads, void* stream, int emulate_ranks) {
  if (!c || ctas < 1 || ctas > HVD_MAX_CTAS || threads > 512 || slot < 0 || slot >= HVD_NUM_SLOTS || ((off | n) & 3)) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  if (emulate_ranks > 0) {
    void* args[] = {(void*)c, (void*)&slot, (void*)&root, (void*)&off, (void*)&n};
    return (int)cudaLaunchCooperativeKernel((void*)hvd_bcast_kernel, dim3(ctas, emulate_ranks), dim3(threads), args, 0, st);
  }
  hvd_bcast_kernel<<<ctas, threads, 0, st>>>(*c, slot, root, off, n);
  return (int)cudaGetLastError();
}

extern "C" int hvd_pad_bytes() { return HVD_NUM_SLOTS * HVD_MAX_CTAS * HVD_MAX_RANKS * (int)sizeof(uint32_t); }
extern "C" int hvd_epoch_bytes() { return HVD_NUM_SLOTS * HVD_MAX_CTAS * (int)sizeof(uint32_t); }

// ------------------------------------------------------------------------------ runtime helpers
extern "C" int hvd_cuda_malloc(void** p, long long bytes) { return (int)cudaMalloc(p, (size_t)bytes); }
extern "C" int hvd_cuda_free(void* p) { return (int)cudaFree(p); }
extern "C" int hvd_ipc_get_handle(void* p, void* out64) { return (int)cudaIpcGetMemHandle((cudaIpcMemHandle_t*)out64, p); }
extern "C" int hvd_ipc_open(const void* h64, void** p) {
  cudaIpcMemHandle_t h;
  memcpy(&h, h64, sizeof(h));
  return (int)cudaIpcOpenMemHandle(p, h, cudaIpcMemLazyEnablePeerAccess);
}
extern "C" int hvd_ipc_close(void* p) { return (int)cudaIpcCloseMemHandle(p); }
extern "C" int hvd_can_access_peer(int dev, int peer) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) return -1;
  return ok;
}
// Plain device-to-device copy on `stream`; src may be a peer-mapped (VMM / IPC) address.
extern "C" int hvd_memcpy_d2d(void* dst, const void* src, long long bytes, void* stream) {
  return (int)cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
}
// Asynchronous copy in any direction on `stream` (the engine's per-step input staging from pinned host
// memory and its 8-byte result read-back: one ctypes call instead of a framework copy under a stream guard).
extern "C" int hvd_memcpy_async(void* dst, const void* src, long long bytes, void* stream) {
  return (int)cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream);
}
// 1 if `p` is page-locked host memory known to the CUDA runtime (safe source for a truly asynchronous H2D copy).
extern "C" int hvd_host_is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
  return a.type == cudaMemoryTypeHost ? 1 : 0;
}
extern "C" const char* hvd_cuda_error_string(int e) { return cudaGetErrorString((cudaError_t)e); }
