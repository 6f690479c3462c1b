// Generic tcgen05 TF32 GEMM:  D[M,N] (fp32, row-major) (+)= A * B,  M tile 128, N tile 64/128, K block 32.
//
// Operands come straight from row-major global tensors through TMA, in either major-ness:
//   A "K-major"  : tensor [M][K]   (rows of K contiguous)      A "MN-major": tensor [K][M]
//   B "K-major"  : tensor [N][K]                                B "MN-major": tensor [K][N]
// so the four combinations cover  X*W, X*W^T, X^T*dY, ...  without any transpose kernels - the
// dense-layer forward (x[B,9216] * W[9216,128]: A K-major, B MN-major), dgrad (dY * W^T: A K-major,
// B K-major) and wgrad (X^T * dY: A MN-major, B MN-major) of the reference's Dense(128) layer
// (mnist_keras.py:79) are all this kernel.
//
// Structure (one CTA per output tile or per split-K slice, 384 threads):
//   warp 0 : TMA producer          (one elected lane; 4-stage smem ring, mbarrier full/empty)
//   warp 1 : MMA issuer            (one elected lane; tcgen05.mma kind::tf32, accumulator in TMEM)
//   warp 2 : TMEM allocator
//   warps 4-11: epilogue           (tcgen05.ld 32x32b -> registers -> fused epilogue -> global; two warps
//                                   per TMEM lane quadrant, each taking every other 32-column chunk)
//
// Two hooks tie it into the pipelined data-parallel step (comm/pipeline.cuh):
//   EPI_RS_PUSH : the epilogue stages the 128 x 128 tile in shared memory and writes it, as whole 512-byte
//                 rows, into the receive area of the rank that owns those weight rows - local memory or a peer
//                 GPU over NVLink - then raises a per-tile flag there (GEMM -> reduce-scatter in one kernel);
//   wait        : the TMA producer warp first waits until the weights it is about to read have been delivered by
//                 their owners (version flags stamped by the all-gather CTAs that run beside the conv forward),
//                 instead of anybody sitting in an exit barrier.
#include "comm/pipeline.cuh"
#include "mnist_head.cuh"
#include "tc_common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 32;          // 32 fp32 = 128 B = one swizzle row
constexpr int STAGES = 4;

// EPI_UNPOOL: D is the conv2 output gradient [B,24,24,64]; the GEMM result (gradient wrt the pooled,
// dropped-out activations, row = sample, column = flat (ph,pw,c)) is routed through the mask byte
// written by conv2_fprop: scaled by 1/keep where live, placed at the arg-max of its 2x2 window,
// zeros at the other three positions - MaxPool/Dropout/ReLU backward fused into the dgrad epilogue.
// EPI_SPLITK_HEAD: split-K accumulation (as EPI_ATOMIC_ADD) followed, in the same kernel, by the classifier head
// (mnist_head.cuh): every slice CTA adds its partial tile, takes a ticket on a device counter, waits until all
// slices of this launch have arrived, and the first ceil(B/8) CTAs then run the head on 8 rows each - the
// Dense(128) forward and everything up to the Dense(128) pre-activation gradient without a kernel boundary.
enum Epilogue { EPI_STORE = 0, EPI_ATOMIC_ADD = 1, EPI_BIAS_RELU = 2, EPI_UNPOOL = 3, EPI_RS_PUSH = 4, EPI_SPLITK_HEAD = 5 };

struct GemmParams {
  int M, N, K;
  int k_split;                   // gridDim.z slices of K
  float* D;
  long long ldd;
  const float* bias;
  int epilogue;
  uint32_t lbo_a, sbo_a, lbo_b, sbo_b;   // 0 = default (test hook for descriptor bring-up)
  const uint8_t* mask;                   // EPI_UNPOOL
  float scale;                           // EPI_UNPOOL: 1/keep
  HvdRsPush rs;                          // EPI_RS_PUSH
  HvdParamWait wait;                     // flags == nullptr: no wait
  mnist::HeadParams head;                // EPI_SPLITK_HEAD
  unsigned long long* head_counter;      // EPI_SPLITK_HEAD: arrival tickets, never reset (one per slice CTA per launch)
};

template <int BN>
struct Smem {
  static constexpr int A_BYTES = BM * BK * 4;     // 16 KB
  static constexpr int B_BYTES = BN * BK * 4;
  static constexpr int STAGE = A_BYTES + B_BYTES;
  static constexpr int TOTAL = STAGES * STAGE + 1024 /*align slack*/ + 256 /*barriers*/;
};

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(384, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Smem<BN>::STAGE);
  uint64_t* full = bars;                 // [STAGES]
  uint64_t* empty = bars + STAGES;       // [STAGES]
  uint64_t* acc_full = bars + 2 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int kb_total = (p.K + BK - 1) / BK;      // a K tail is zero-filled by TMA
  const int kb_per = (kb_total + p.k_split - 1) / p.k_split;
  const int kb_begin = blockIdx.z * kb_per;
  const int kb_end = min(kb_total, kb_begin + kb_per);
  const int nkb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], 1); }
    tc::mbar_init(acc_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc<BN>(tmem_slot);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  tc::pdl_launch_dependents();
  tc::pdl_wait();                       // everything above overlapped the previous kernel's tail

  if (warp == 0) {
    hvd_param_wait_warp(p.wait);        // all 32 lanes poll the delivery flags in parallel
    if (tc::elect_one()) {
      for (int i = 0; i < nkb; i++) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        tc::mbar_wait(&empty[s], ph ^ 1);
        uint8_t* a = smem + s * Smem<BN>::STAGE;
        uint8_t* b = a + Smem<BN>::A_BYTES;
        tc::mbar_expect_tx(&full[s], Smem<BN>::STAGE);
        const int k0 = (kb_begin + i) * BK;
        if (!A_MN) tc::tma_load_2d(a, &tmA, &full[s], k0, m0);
        else
          for (int j = 0; j < BM / 32; j++) tc::tma_load_2d(a + j * (BK * 128), &tmA, &full[s], m0 + 32 * j, k0);
        if (!B_MN) tc::tma_load_2d(b, &tmB, &full[s], k0, n0);
        else
          for (int j = 0; j < BN / 32; j++) tc::tma_load_2d(b + j * (BK * 128), &tmB, &full[s], n0 + 32 * j, k0);
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      const uint32_t lbo_a = p.lbo_a ? p.lbo_a : (A_MN ? BK * 128 : 16), sbo_a = p.sbo_a ? p.sbo_a : (A_MN ? 512 : 1024);
      const uint32_t lbo_b = p.lbo_b ? p.lbo_b : (B_MN ? BK * 128 : 16), sbo_b = p.sbo_b ? p.sbo_b : (B_MN ? 512 : 1024);
      for (int i = 0; i < nkb; i++) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        tc::mbar_wait(&full[s], ph);
        tc::tc_fence_after();
        const uint32_t a = tc::smem_u32(smem + s * Smem<BN>::STAGE);
        const uint32_t b = a + Smem<BN>::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 8; k++) {        // UMMA_K = 8 for tf32
          const uint64_t da = tc::make_smem_desc(a + (A_MN ? k * 1024 : k * 32), lbo_a, sbo_a, A_MN ? 1 : 2);
          const uint64_t db = tc::make_smem_desc(b + (B_MN ? k * 1024 : k * 32), lbo_b, sbo_b, B_MN ? 1 : 2);
          tc::mma_tf32(tmem, da, db, idesc, (i | k) != 0);
        }
        tc::mma_commit(&empty[s]);                // frees the smem stage when these MMAs retire
      }
      tc::mma_commit(acc_full);
    }
  } else if (warp >= 4) {
    const int q = warp & 3;                       // TMEM lane quadrant this warp may read
    tc::mbar_wait(acc_full, 0);
    tc::tc_fence_after();
    const int row = m0 + q * 32 + lane;
#pragma unroll 1
    for (int c = ((warp - 4) >> 2) * 32; c < BN; c += 64) {     // two warps per lane quadrant split the column chunks
      float v[32];
      tc::tmem_ld_x32(tmem + ((uint32_t)(q * 32) << 16) + c, v);
      tc::tmem_ld_wait();
      if (nkb <= 0) {
#pragma unroll
        for (int j = 0; j < 32; j++) v[j] = 0.f;
      }
      if (p.epilogue == EPI_UNPOOL) {
        // The accumulator gives each lane ONE sample's 32 channels, but consecutive samples are 147 KB apart
        // in dY2: stored from that layout every warp store touches 32 half-filled sectors.  Transpose through
        // shared memory (the pipeline stages are idle once the accumulator is complete) so that 8 lanes
        // write one sample's 128 contiguous bytes and a warp store fills 16 whole sectors.
        float* T = reinterpret_cast<float*>(smem + (warp - 4) * (32 * 33 * 4 + 32 * 9 * 4));   // [32][33] values
        uint32_t* Mw = reinterpret_cast<uint32_t*>(T + 32 * 33);                                // [32][9] mask words
        const int f = n0 + c;                                  // flat (ph, pw, c0) with c0 in {0, 32}
        const int ph = f / (12 * 64), pw = (f / 64) % 12, c0 = f & 63;
        {
          uint4 m0 = make_uint4(0, 0, 0, 0), m1 = m0;          // rows past M: nothing live
          if (row < p.M) {
            const uint4* mp = reinterpret_cast<const uint4*>(p.mask + (long long)row * 9216 + f);
            m0 = mp[0]; m1 = mp[1];
          }
          const uint32_t mw[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
          for (int j = 0; j < 32; j++) T[lane * 33 + j] = v[j] * p.scale;
#pragma unroll
          for (int j = 0; j < 8; j++) Mw[lane * 9 + j] = mw[j];
        }
        __syncwarp();
        const int k = lane & 7;                                // channels 4k .. 4k+3 of the chunk
#pragma unroll 2
        for (int i = 0; i < 8; i++) {
          const int b = i * 4 + (lane >> 3);                   // sample inside this warp's 32 rows
          const int grow = m0 + q * 32 + b;
          const float t0 = T[b * 33 + 4 * k], t1 = T[b * 33 + 4 * k + 1], t2 = T[b * 33 + 4 * k + 2], t3 = T[b * 33 + 4 * k + 3];
          const uint32_t mk = Mw[b * 9 + k];
          if (grow < p.M) {
            float* base = p.D + (((long long)grow * 24 + 2 * ph) * 24 + 2 * pw) * 64 + c0 + 4 * k;
            const float tv[4] = {t0, t1, t2, t3};
#pragma unroll
            for (int pos = 0; pos < 4; pos++) {
              float o[4];
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const uint32_t mb = (mk >> (8 * e)) & 0xff;
                o[e] = ((mb & 4) && (int)(mb & 3) == pos) ? tv[e] : 0.f;
              }
              *reinterpret_cast<float4*>(base + ((pos >> 1) * 24 + (pos & 1)) * 64) = make_float4(o[0], o[1], o[2], o[3]);
            }
          }
        }
        __syncwarp();                                          // the next column chunk reuses T / Mw
      } else if (p.epilogue == EPI_RS_PUSH) {
        // stage the tile row-major in the (now idle) operand ring, 16-byte chunks XOR-swizzled by the row so
        // that neither these per-row writes nor the per-row reads below conflict on banks
        float4* T = reinterpret_cast<float4*>(smem);
        const int rl = q * 32 + lane;
#pragma unroll
        for (int j = 0; j < 8; j++)
          T[rl * 32 + (((c >> 2) + j) ^ lane)] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      } else if (row < p.M) {
        float* d = p.D + (long long)row * p.ldd + n0 + c;
        if (p.epilogue == EPI_ATOMIC_ADD || p.epilogue == EPI_SPLITK_HEAD) {
          if (n0 + c + 32 <= p.N && (p.ldd & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) tc::red_add_f4(d + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
            for (int j = 0; j < 32; j++) if (n0 + c + j < p.N) atomicAdd(d + j, v[j]);
          }
        } else {
          if (p.epilogue == EPI_BIAS_RELU) {
#pragma unroll
            for (int j = 0; j < 32; j++) v[j] = fmaxf(v[j] + p.bias[n0 + c + j], 0.f);
          }
          if (n0 + c + 32 <= p.N) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(d + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
            for (int j = 0; j < 32; j++) if (n0 + c + j < p.N) d[j] = v[j];
          }
        }
      }
    }
  }
  if (warp >= 4 && p.epilogue == EPI_RS_PUSH) {
    // all 8 epilogue warps: tile staged -> every warp store is one whole 512-byte row of the destination
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const HvdRsPush& rs = p.rs;
    const int tile = blockIdx.x;
    const int owner = tile / rs.tiles_per_owner, tl = tile - owner * rs.tiles_per_owner;
    float* dst = rs.recv[owner] + (long long)rs.rank * rs.shard_elems + (long long)tl * (BM * 128);
    const float4* T = reinterpret_cast<const float4*>(smem);
    const int w = warp - 4;
#pragma unroll 4
    for (int r = w * 16; r < w * 16 + 16; r++)
      reinterpret_cast<float4*>(dst + r * 128)[lane] = T[r * 32 + (lane ^ (r & 31))];
    if (rs.nranks > 1 && rs.send_flags) {
      // every epilogue thread's rows happen-before the barrier; the release below is cumulative over them, so
      // one system fence (inside st.release.sys) is enough - no fence per thread
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 128) {                    // the owner is told that the tile has landed
        uint32_t* ep = rs.epoch + HVD_SLOT_RS_PUSH * HVD_MAX_CTAS + tile;
        const uint32_t e = *ep + 1;
        if (owner != rs.rank)
          st_release_sys(rs.pad[owner] + ((size_t)(HVD_SLOT_RS_PUSH * HVD_MAX_CTAS + tl) * HVD_MAX_RANKS + rs.rank), e);
        *ep = e;
      }
    }
  }
  if (warp >= 4 && p.epilogue == EPI_SPLITK_HEAD) {
    // all 8 epilogue warps.  This CTA's partial sums are in L2 (red.add) before its ticket is taken ...
    __threadfence();
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x == 128) {
      __threadfence();
      const unsigned long long ks = gridDim.z;
      const unsigned long long old = atomicAdd(p.head_counter, 1ull);
      const unsigned long long target = (old / ks + 1ull) * ks;       // all slices of THIS launch
      unsigned spins = 0;
      for (;;) {
        unsigned long long now;
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(now) : "l"(p.head_counter) : "memory");
        if (now >= target) break;
        __nanosleep(32);
        if (++spins > (1u << 24)) { asm volatile("trap;"); }           // a lost slice must trap, not hang the GPU
      }
      __threadfence();
    }
    // ... and every slice's sums are there before anybody reads them (the slices are co-resident: k_split <= SMs)
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if ((int)blockIdx.z * mnist::HEAD_ROWS < p.head.B) {
      mnist::HeadSmem& sm = *reinterpret_cast<mnist::HeadSmem*>(smem);    // the operand ring is idle by now
      mnist::head_rows(p.head, blockIdx.z, threadIdx.x - 128, sm, [] { asm volatile("bar.sync 1, 256;" ::: "memory"); });
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc<BN>(tmem);
}

template <int BN, bool A_MN, bool B_MN>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st) {
  auto k = tc_gemm_kernel<BN, A_MN, B_MN>;
  static tc::PerDeviceOnce attr;
  if (attr.need()) {
    if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<BN>::TOTAL) != cudaSuccess) return -10;
    attr.mark();
  }
  dim3 grid((p.M + BM - 1) / BM, (p.N + BN - 1) / BN, p.k_split);
  return (int)tc::launch_pdl(k, grid, dim3(384), Smem<BN>::TOTAL, st, ta, tb, p);
}

}  // namespace

// A: [M][K] (a_mn=0) or [K][M] (a_mn=1); B: [N][K] (b_mn=0) or [K][N] (b_mn=1); all dense row-major fp32.
// M/N/K tails are zero-filled by TMA (out-of-bounds box elements) and masked in the epilogue.
extern "C" int hvd_tc_gemm_ex(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn, int bn,
                              int k_split, int epilogue, const float* bias, long long ldd, void* stream,
                              unsigned lbo_a, unsigned sbo_a, unsigned lbo_b, unsigned sbo_b, const uint8_t* mask,
                              float scale, const HvdRsPush* rs, const HvdParamWait* wait) {
  if ((bn != 64 && bn != 128) || k_split < 1 || K < 1) return -1;
  if (epilogue == EPI_RS_PUSH) {
    // whole 128 x 128 tiles of a [M][128] result, one owner per run of tiles
    if (!rs || bn != 128 || N != 128 || (M % BM) != 0 || k_split != 1 || rs->nranks < 1 || rs->nranks > HVD_MAX_RANKS ||
        rs->tiles_per_owner < 1 || (M / BM) != rs->tiles_per_owner * rs->nranks || (M / BM) > HVD_MAX_CTAS)
      return -5;
  }
  CUtensorMap ta, tb;
  {
    uint64_t dims[2], str[1];
    uint32_t box[2];
    if (!a_mn) { dims[0] = K; dims[1] = M; str[0] = (uint64_t)K * 4; box[0] = BK; box[1] = BM; }
    else { dims[0] = M; dims[1] = K; str[0] = (uint64_t)M * 4; box[0] = 32; box[1] = BK; }
    if (tc::make_tmap(&ta, A, 2, dims, str, box, a_mn != 0)) return -2;
    if (!b_mn) { dims[0] = K; dims[1] = N; str[0] = (uint64_t)K * 4; box[0] = BK; box[1] = (uint32_t)bn; }
    else { dims[0] = N; dims[1] = K; str[0] = (uint64_t)N * 4; box[0] = 32; box[1] = BK; }
    if (tc::make_tmap(&tb, B, 2, dims, str, box, b_mn != 0)) return -3;
  }
  if (epilogue == EPI_SPLITK_HEAD) return -6;          // only through hvd_dense1_fwd_head
  GemmParams p{M, N, K, k_split, D, ldd ? ldd : (long long)N, bias, epilogue, lbo_a, sbo_a, lbo_b, sbo_b, mask, scale, {}, {},
               {}, nullptr};
  if (rs) p.rs = *rs;
  if (wait) p.wait = *wait;
  cudaStream_t st = (cudaStream_t)stream;
#define HVD_CASE(BN_, AM, BMN) if (bn == BN_ && a_mn == AM && b_mn == BMN) return launch<BN_, AM != 0, BMN != 0>(ta, tb, p, st);
  HVD_CASE(64, 0, 0) HVD_CASE(64, 0, 1) HVD_CASE(64, 1, 0) HVD_CASE(64, 1, 1)
  HVD_CASE(128, 0, 0) HVD_CASE(128, 0, 1) HVD_CASE(128, 1, 0) HVD_CASE(128, 1, 1)
#undef HVD_CASE
  return -4;
}

extern "C" int hvd_tc_gemm(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn, int bn,
                           int k_split, int epilogue, const float* bias, long long ldd, void* stream,
                           unsigned lbo_a, unsigned sbo_a, unsigned lbo_b, unsigned sbo_b, const uint8_t* mask,
                           float scale) {
  return hvd_tc_gemm_ex(A, B, D, M, N, K, a_mn, b_mn, bn, k_split, epilogue, bias, ldd, stream, lbo_a, sbo_a, lbo_b,
                        sbo_b, mask, scale, nullptr, nullptr);
}

// Dense(128) forward + classifier head in ONE launch (reference layers mnist_keras.py:79-81, loss/metric :89-91):
//   h_pre[B][128] += x[B][K] * W[K][128] over k_split slices of K, then hvd_head's work on the reduced sums
// (arguments as hvd_head; h_pre must be zero on entry and is zero again on exit).  `counter` is a device
// uint64 that starts at 0 and is never reset.  Needs k_split * 8 >= B (one head group of 8 rows per slice CTA)
// and k_split <= 148 (the slices wait for each other).  wait: as hvd_tc_gemm_ex (may be NULL).
extern "C" int hvd_dense1_fwd_head(const float* x, const float* W, int B, int K, int k_split, float* h_pre, const float* b1,
                                   const float* w2, const float* b2, const long long* labels, float* dh_pre, float* g_w2,
                                   float* g_b2, float* g_b1, float* probs, void* step_state, int training, int backward,
                                   float keep, unsigned long long* counter, const HvdParamWait* wait, void* stream) {
  if (B < 1 || B > BM || K < 1 || (K & 3) || k_split < 1 || k_split > 148 || k_split * mnist::HEAD_ROWS < B || !counter) return -1;
  CUtensorMap ta, tb;
  {
    uint64_t dims[2], str[1];
    uint32_t box[2];
    dims[0] = K; dims[1] = B; str[0] = (uint64_t)K * 4; box[0] = BK; box[1] = BM;
    if (tc::make_tmap(&ta, x, 2, dims, str, box, false)) return -2;
    dims[0] = 128; dims[1] = K; str[0] = 128 * 4; box[0] = 32; box[1] = BK;
    if (tc::make_tmap(&tb, W, 2, dims, str, box, true)) return -3;
  }
  GemmParams p{B, 128, K, k_split, h_pre, 128, nullptr, EPI_SPLITK_HEAD, 0, 0, 0, 0, nullptr, 1.0f, {}, {}, {}, counter};
  if (wait) p.wait = *wait;
  p.head = mnist::HeadParams{B, training, backward, mnist::HEAD_ROWS, keep, h_pre, b1, w2, b2, labels, dh_pre, g_w2, g_b2,
                             g_b1, probs, reinterpret_cast<mnist::StepState*>(step_state)};
  return launch<128, false, true>(ta, tb, p, (cudaStream_t)stream);
}
