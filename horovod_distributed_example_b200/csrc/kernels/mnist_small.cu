// CUDA-core kernels of the fused MNIST engine: the ops whose GEMM shape is degenerate
// (conv1: K = 9; dense2: N = 10) or that are pure reductions.  Tensor cores would idle on
// these; they are bandwidth/latency bound and run at memory speed here.
//   conv1_fwd   : Conv2D(32, 3x3, relu) on the 1-channel input            (mnist_keras.py:72-74)
//   head        : bias+relu+dropout(0.5) of dense1, Dense(10)+softmax, cross-entropy, accuracy, and
//                 the whole backward of dense2 + the dropout/relu backward of dense1 in one CTA
//                                                                          (mnist_keras.py:79-81,89-91)
//   conv1_wgrad : dW/db of conv1 and db of conv2 (pure reductions)
#include "comm/pipeline.cuh"
#include "mnist_common.cuh"
#include "mnist_head.cuh"
#include "tc_common.cuh"

namespace {
using namespace mnist;

// ------------------------------------------------------------------------------- conv1 forward
// One thread = one output pixel x 8 channels; the 9 taps come from L1, the 288 weights from smem.
// It is also the step prologue: it zeroes the gradient ranges later kernels accumulate into with
// red.add, clears the per-step loss/accuracy words and advances the dropout step counter, so the
// captured graph needs no memset nodes.
struct Conv1Prologue {
  float* zero_a; int n_a;       // float4-aligned ranges to clear (may be null / 0)
  float* zero_b; int n_b;
  StepState* st;                // may be null
  int advance_step;
};

// Pipelined data-parallel step: the first `ag_blocks` blocks of this grid do not convolve - they deliver the rows
// of dense_1/kernel this rank updated in the previous step to the other replicas (hvd_allgather_block), hidden
// under the convolution the remaining blocks compute.
struct Conv1Comm {
  HvdCommCtx c;
  HvdAllgatherArgs ag;
  int ag_blocks;
};

__global__ void __launch_bounds__(256, 5) conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w /*[9][32]*/,
                                                        const float* __restrict__ bias, float* __restrict__ a1, int B,
                                                        Conv1Prologue pro, const __grid_constant__ Conv1Comm comm) {
  __shared__ float sw[9 * 32 + 32];
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  if ((int)blockIdx.x < comm.ag_blocks) {
    hvd_allgather_block(comm.c, comm.ag, reinterpret_cast<float**>(sw), blockIdx.x, comm.ag_blocks);
    return;
  }
  const int cblk = blockIdx.x - comm.ag_blocks, cgrid = gridDim.x - comm.ag_blocks;
  {
    const int gid = cblk * blockDim.x + threadIdx.x, gsz = cgrid * blockDim.x;
    for (int i = gid; i < pro.n_a / 4; i += gsz) reinterpret_cast<float4*>(pro.zero_a)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = gid; i < pro.n_b / 4; i += gsz) reinterpret_cast<float4*>(pro.zero_b)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (gid == 0 && pro.st) {
      pro.st->last_loss = 0.f;
      pro.st->last_correct = 0.f;
      if (pro.advance_step) pro.st->step += 1;
    }
  }
  for (int i = threadIdx.x; i < 9 * 32 + 32; i += blockDim.x) sw[i] = i < 288 ? w[i] : bias[i - 288];
  __syncthreads();
  // Two work items per thread, both items' loads issued before either is used, and a grid of ONE wave (676 CTAs
  // at batch 128, five resident per SM): with one item per thread the 346 K items needed a partial second round
  // on the 303 K resident threads, i.e. two dependent load-compute-store latencies instead of one.
  const long long total = (long long)B * H1 * H1 * 4;
  const long long half = (total + 1) / 2;
  for (long long t0 = (long long)cblk * blockDim.x + threadIdx.x; t0 < half; t0 += (long long)cgrid * blockDim.x) {
    float in[2][9];
    long long pixs[2];
    int cgs[2];
    bool on[2];
#pragma unroll
    for (int u = 0; u < 2; u++) {
      const long long t = t0 + u * half;
      on[u] = t < total;
      const long long tt = on[u] ? t : 0;
      cgs[u] = (int)(tt & 3);
      pixs[u] = tt >> 2;
      const int ox = (int)(pixs[u] % H1), oy = (int)((pixs[u] / H1) % H1);
      const int b = (int)(pixs[u] / (H1 * H1));
      const float* xi = x + ((long long)b * IMG + oy) * IMG + ox;
#pragma unroll
      for (int kh = 0; kh < 3; kh++)
#pragma unroll
        for (int kw = 0; kw < 3; kw++) in[u][kh * 3 + kw] = __ldg(xi + kh * IMG + kw);
    }
#pragma unroll
    for (int u = 0; u < 2; u++) {
      if (!on[u]) continue;
      float o[8];
#pragma unroll
      for (int c = 0; c < 8; c++) {
        float acc = sw[288 + cgs[u] * 8 + c];
#pragma unroll
        for (int k = 0; k < 9; k++) acc = fmaf(in[u][k], sw[k * 32 + cgs[u] * 8 + c], acc);
        o[c] = fmaxf(acc, 0.f);
      }
      float4* dst = reinterpret_cast<float4*>(a1 + pixs[u] * C1 + cgs[u] * 8);
      dst[0] = make_float4(o[0], o[1], o[2], o[3]);
      dst[1] = make_float4(o[4], o[5], o[6], o[7]);
    }
  }
}

// ------------------------------------------------------------------------------- head
// (body in mnist_head.cuh: the same code runs inside the split-K Dense(128) forward GEMM, tc_gemm.cu)
// Grid of ceil(B/8) CTAs x 256 threads; warp w owns batch row 8*cta + w.
__global__ void __launch_bounds__(256) head_kernel(HeadParams p) {
  __shared__ HeadSmem sm;
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  head_rows(p, blockIdx.x, threadIdx.x, sm, [] { __syncthreads(); });
}

// ------------------------------------------------------------------------------- conv1 wgrad (+ bias grads)
// lane = channel; each warp walks pixels, keeps 9 tap sums + 1 bias sum per lane; block -> atomics.
__global__ void __launch_bounds__(256) conv1_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ da1,
                                                          float* g_w /*[9][32]*/, float* g_b /*[32]*/, int B) {
  __shared__ float red[8][10][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long npix = (long long)B * H1 * H1;
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  float acc[10];
#pragma unroll
  for (int k = 0; k < 10; k++) acc[k] = 0.f;
  constexpr int U = 4;                      // pixels in flight per warp (memory-level parallelism)
  for (long long pix0 = ((long long)blockIdx.x * 8 + warp) * U; pix0 < npix; pix0 += (long long)gridDim.x * 8 * U) {
    float g[U], xv[U][9];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const long long pix = pix0 + u < npix ? pix0 + u : npix - 1;
      const int ox = (int)(pix % H1), oy = (int)((pix / H1) % H1);
      const int b = (int)(pix / (H1 * H1));
      g[u] = pix0 + u < npix ? __ldg(da1 + pix * C1 + lane) : 0.f;
      const float* xi = x + ((long long)b * IMG + oy) * IMG + ox;
#pragma unroll
      for (int kh = 0; kh < 3; kh++)
#pragma unroll
        for (int kw = 0; kw < 3; kw++) xv[u][kh * 3 + kw] = __ldg(xi + kh * IMG + kw);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
#pragma unroll
      for (int k = 0; k < 9; k++) acc[k] = fmaf(xv[u][k], g[u], acc[k]);
      acc[9] += g[u];
    }
  }
#pragma unroll
  for (int k = 0; k < 10; k++) red[warp][k][lane] = acc[k];
  __syncthreads();
  for (int o = threadIdx.x; o < 320; o += blockDim.x) {
    const int k = o / 32, c = o % 32;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; w++) s += red[w][k][c];
    if (k < 9) atomicAdd(g_w + k * 32 + c, s);
    else atomicAdd(g_b + c, s);
  }
}

// column sums of a [rows][64] matrix -> atomics (conv2 bias gradient = sum over pixels of dY2)
__global__ void __launch_bounds__(256) colsum64_kernel(const float* __restrict__ m, float* out, long long rows) {
  __shared__ float red[4][64];
  const int c = threadIdx.x & 63, g = threadIdx.x >> 6;
  tc::pdl_launch_dependents();
  tc::pdl_wait();
  float acc = 0.f;
  for (long long r = (long long)blockIdx.x * 4 + g; r < rows; r += (long long)gridDim.x * 4) acc += m[r * 64 + c];
  red[g][c] = acc;
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(out + c, red[0][c] + red[1][c] + red[2][c] + red[3][c]);
}

}  // namespace

// blocks_per_sm (1..8): resident 256-thread CTAs per SM the grid is sized for.  8 fills every thread slot; the
// pipelined multi-GPU step passes 7 so that the small all-gather CTAs enqueued beside it find a slot at once.
// comm / ag / ag_blocks (may be NULL / 0): the all-gather role of the first ag_blocks blocks, see Conv1Comm.
extern "C" int hvd_conv1_fwd_ex(const float* x, const float* w, const float* bias, float* a1, int B, float* zero_a,
                                int n_a, float* zero_b, int n_b, void* step_state, int advance_step, void* stream,
                                int blocks_per_sm, const HvdCommCtx* comm, const HvdAllgatherArgs* ag, int ag_blocks) {
  if ((n_a | n_b) & 3) return -1;
  if (blocks_per_sm < 1 || blocks_per_sm > 8) return -2;
  const long long total = (long long)B * mnist::H1 * mnist::H1 * 4;
  int blocks = (int)(((total + 1) / 2 + 255) / 256);          // two work items per thread
  if (blocks > 148 * blocks_per_sm) blocks = 148 * blocks_per_sm;
  static Conv1Comm none;               // zero-initialised: ag_blocks = 0
  Conv1Comm cc = none;
  if (comm && ag && ag_blocks > 0) {
    if (comm->nranks < 2 || ag_blocks > HVD_MAX_CTAS || (ag->use_mc && comm->mc_base == nullptr)) return -3;
    cc.c = *comm; cc.ag = *ag; cc.ag_blocks = ag_blocks;
    blocks += ag_blocks;
  }
  Conv1Prologue pro{zero_a, zero_a ? n_a : 0, zero_b, zero_b ? n_b : 0, reinterpret_cast<mnist::StepState*>(step_state), advance_step};
  return (int)tc::launch_pdl(conv1_fwd_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, x, w, bias, a1, B, pro, cc);
}

extern "C" int hvd_conv1_fwd(const float* x, const float* w, const float* bias, float* a1, int B, float* zero_a,
                             int n_a, float* zero_b, int n_b, void* step_state, int advance_step, void* stream) {
  return hvd_conv1_fwd_ex(x, w, bias, a1, B, zero_a, n_a, zero_b, n_b, step_state, advance_step, stream, 8, nullptr,
                          nullptr, 0);
}

extern "C" int hvd_head(float* h_pre, const float* b1, const float* w2, const float* b2, const long long* labels,
                        float* dh_pre, float* g_w2, float* g_b2, float* g_b1, float* probs, void* step_state, int B,
                        int training, int backward, float keep, void* stream) {
  if (B < 1) return -1;
  HeadParams p{B, training, backward, HEAD_ROWS, keep, h_pre, b1, w2, b2, labels, dh_pre, g_w2, g_b2, g_b1, probs,
               reinterpret_cast<mnist::StepState*>(step_state)};
  return (int)tc::launch_pdl(head_kernel, dim3((B + HEAD_ROWS - 1) / HEAD_ROWS), dim3(256), 0, (cudaStream_t)stream, p);
}

extern "C" int hvd_conv1_wgrad(const float* x, const float* da1, float* g_w, float* g_b, int B, void* stream) {
  return (int)tc::launch_pdl(conv1_wgrad_kernel, dim3(148 * 4), dim3(256), 0, (cudaStream_t)stream, x, da1, g_w, g_b, B);
}

extern "C" int hvd_colsum64(const float* m, float* out, long long rows, void* stream) {
  colsum64_kernel<<<148 * 2, 256, 0, (cudaStream_t)stream>>>(m, out, rows);
  return (int)cudaGetLastError();
}
