// CUDA-core kernels of the fused MNIST engine: the ops whose GEMM shape is degenerate
// (conv1: K = 9; dense2: N = 10) or that are pure reductions.  Tensor cores would idle on
// these; they are bandwidth/latency bound and run at memory speed here.
//   conv1_fwd   : Conv2D(32, 3x3, relu) on the 1-channel input            (mnist_keras.py:72-74)
//   head        : bias+relu+dropout(0.5) of dense1, Dense(10)+softmax, cross-entropy, accuracy, and
//                 the whole backward of dense2 + the dropout/relu backward of dense1 in one CTA
//                                                                          (mnist_keras.py:79-81,89-91)
//   conv1_wgrad : dW/db of conv1 and db of conv2 (pure reductions)
#include "mnist_common.cuh"

namespace {
using namespace mnist;

// ------------------------------------------------------------------------------- conv1 forward
// One thread = one output pixel x 8 channels; the 9 taps come from L1, the 288 weights from smem.
__global__ void __launch_bounds__(256) conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w /*[9][32]*/,
                                                        const float* __restrict__ bias, float* __restrict__ a1, int B) {
  __shared__ float sw[9 * 32 + 32];
  for (int i = threadIdx.x; i < 9 * 32 + 32; i += blockDim.x) sw[i] = i < 288 ? w[i] : bias[i - 288];
  __syncthreads();
  const long long total = (long long)B * H1 * H1 * 4;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(t & 3);
    const long long pix = t >> 2;
    const int ox = (int)(pix % H1), oy = (int)((pix / H1) % H1);
    const int b = (int)(pix / (H1 * H1));
    const float* xi = x + ((long long)b * IMG + oy) * IMG + ox;
    float in[9];
#pragma unroll
    for (int kh = 0; kh < 3; kh++)
#pragma unroll
      for (int kw = 0; kw < 3; kw++) in[kh * 3 + kw] = __ldg(xi + kh * IMG + kw);
    float o[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
      float acc = sw[288 + cg * 8 + c];
#pragma unroll
      for (int k = 0; k < 9; k++) acc = fmaf(in[k], sw[k * 32 + cg * 8 + c], acc);
      o[c] = fmaxf(acc, 0.f);
    }
    float4* dst = reinterpret_cast<float4*>(a1 + pix * C1 + cg * 8);
    dst[0] = make_float4(o[0], o[1], o[2], o[3]);
    dst[1] = make_float4(o[4], o[5], o[6], o[7]);
  }
}

// ------------------------------------------------------------------------------- head
struct HeadParams {
  int B, training, backward;
  float keep;
  float* h_pre;               // [B,128] split-K sums of dense1 (zeroed again after use)
  const float* b1;            // dense1 bias [128]
  const float* w2;            // dense2 kernel [128,10]
  const float* b2;            // [10]
  const long long* labels;    // [B] int64 class ids
  float* dh_pre;              // [B,128] out: gradient wrt dense1 pre-activation
  float* g_w2; float* g_b2; float* g_b1;   // gradient slots in the arena
  float* probs;               // [B,10] out (predict / eval), may be null
  StepState* st;
};

// Single CTA, 1024 threads; everything lives in shared memory (h: 64 KB).
__global__ void __launch_bounds__(1024, 1) head_kernel(HeadParams p) {
  extern __shared__ float sm[];
  float* h = sm;                        // [B][128] post relu+dropout activations
  float* msk = h + p.B * HID;           // [B][128] d(h)/d(pre): 0 or 1/keep
  float* w2 = msk + p.B * HID;          // [128][10]
  float* dl = w2 + HID * NCLS;          // [B][10] logits -> probs -> dlogits
  float* red = dl + p.B * NCLS;         // [64] scratch
  const int tid = threadIdx.x, nt = blockDim.x;
  const unsigned step = p.st->step, seed = p.st->seed;
  const float inv_keep = 1.0f / p.keep;

  for (int i = tid; i < HID * NCLS; i += nt) w2[i] = p.w2[i];
  for (int i4 = tid; i4 < p.B * HID / 4; i4 += nt) {
    float4 v = reinterpret_cast<float4*>(p.h_pre)[i4];
    reinterpret_cast<float4*>(p.h_pre)[i4] = make_float4(0.f, 0.f, 0.f, 0.f);   // re-arm the split-K accumulator
    const int j = (i4 * 4) & (HID - 1);
    uint4 rnd = make_uint4(0, 0, 0, 0);
    if (p.training) rnd = philox4x32_10(make_uint4((unsigned)i4, step, 0x1u, 0u), make_uint2(seed, 0x5EEDu));
    const float pre[4] = {v.x + p.b1[j], v.y + p.b1[j + 1], v.z + p.b1[j + 2], v.w + p.b1[j + 3]};
    const unsigned rr[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const bool keep = !p.training || u01(rr[e]) < p.keep;
      const float scale = p.training ? inv_keep : 1.0f;
      const bool live = keep && pre[e] > 0.f;
      h[i4 * 4 + e] = live ? pre[e] * scale : 0.f;
      msk[i4 * 4 + e] = live ? scale : 0.f;
    }
  }
  __syncthreads();
  // logits[b][k]
  for (int o = tid; o < p.B * NCLS; o += nt) {
    const int b = o / NCLS, k = o % NCLS;
    float acc = p.b2[k];
    const float* hb = h + b * HID;
#pragma unroll 8
    for (int j = 0; j < HID; j++) acc = fmaf(hb[j], w2[j * NCLS + k], acc);
    dl[o] = acc;
  }
  __syncthreads();
  // softmax, loss, accuracy, dlogits: one thread per sample
  float my_loss = 0.f, my_ok = 0.f;
  if (tid < p.B) {
    float* l = dl + tid * NCLS;
    float mx = l[0];
    int am = 0;
#pragma unroll
    for (int k = 1; k < NCLS; k++) if (l[k] > mx) { mx = l[k]; am = k; }
    float e[NCLS], s = 0.f;
#pragma unroll
    for (int k = 0; k < NCLS; k++) { e[k] = expf(l[k] - mx); s += e[k]; }
    const int y = (int)p.labels[tid];
    const float inv = 1.0f / s;
    const float py = fminf(fmaxf(e[y] * inv, 1e-7f), 1.0f - 1e-7f);     // Keras clips probabilities
    my_loss = -logf(py);
    my_ok = (am == y) ? 1.f : 0.f;
    const float invB = 1.0f / (float)p.B;
#pragma unroll
    for (int k = 0; k < NCLS; k++) {
      const float pk = e[k] * inv;
      if (p.probs) p.probs[tid * NCLS + k] = pk;
      l[k] = (pk - (k == y ? 1.f : 0.f)) * invB;
    }
  }
  // block reduction of loss / correct
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    my_loss += __shfl_xor_sync(0xffffffffu, my_loss, o);
    my_ok += __shfl_xor_sync(0xffffffffu, my_ok, o);
  }
  if ((tid & 31) == 0) { red[tid >> 5] = my_loss; red[32 + (tid >> 5)] = my_ok; }
  __syncthreads();
  if (tid == 0) {
    float ls = 0.f, ok = 0.f;
    for (int i = 0; i < nt / 32; i++) { ls += red[i]; ok += red[32 + i]; }
    const float mean = ls / (float)p.B;
    p.st->last_loss = p.backward ? mean : ls;      // eval mode reports sums (weighted by batch size)
    p.st->last_correct = ok;
    p.st->loss_sum += mean;
    p.st->correct_sum += ok;
    if (p.backward) p.st->step = step + 1;
  }
  if (!p.backward) return;
  // dW2[j][k] = sum_b h[b][j] dl[b][k];  db2[k] = sum_b dl[b][k]
  for (int o = tid; o < HID * NCLS + NCLS; o += nt) {
    float acc = 0.f;
    if (o < HID * NCLS) {
      const int j = o / NCLS, k = o % NCLS;
#pragma unroll 8
      for (int b = 0; b < p.B; b++) acc = fmaf(h[b * HID + j], dl[b * NCLS + k], acc);
      p.g_w2[o] = acc;
    } else {
      const int k = o - HID * NCLS;
#pragma unroll 8
      for (int b = 0; b < p.B; b++) acc += dl[b * NCLS + k];
      p.g_b2[k] = acc;
    }
  }
  __syncthreads();                       // dW2 above still reads h; the loop below overwrites it
  // dh_pre[b][j] = (sum_k dl[b][k] w2[j][k]) * mask;   db1[j] = sum_b dh_pre[b][j]
  for (int o = tid; o < p.B * HID; o += nt) {
    const int b = o / HID, j = o % HID;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < NCLS; k++) acc = fmaf(dl[b * NCLS + k], w2[j * NCLS + k], acc);
    acc *= msk[o];
    p.dh_pre[o] = acc;
    h[o] = acc;                          // reuse h for the column sums below
  }
  __syncthreads();
  for (int j = tid; j < HID; j += nt) {
    float acc = 0.f;
#pragma unroll 8
    for (int b = 0; b < p.B; b++) acc += h[b * HID + j];
    p.g_b1[j] = acc;
  }
}

// ------------------------------------------------------------------------------- conv1 wgrad (+ bias grads)
// lane = channel; each warp walks pixels, keeps 9 tap sums + 1 bias sum per lane; block -> atomics.
__global__ void __launch_bounds__(256) conv1_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ da1,
                                                          float* g_w /*[9][32]*/, float* g_b /*[32]*/, int B) {
  __shared__ float red[8][10][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long npix = (long long)B * H1 * H1;
  float acc[10];
#pragma unroll
  for (int k = 0; k < 10; k++) acc[k] = 0.f;
  for (long long pix = (long long)blockIdx.x * 8 + warp; pix < npix; pix += (long long)gridDim.x * 8) {
    const int ox = (int)(pix % H1), oy = (int)((pix / H1) % H1);
    const int b = (int)(pix / (H1 * H1));
    const float g = da1[pix * C1 + lane];
    const float* xi = x + ((long long)b * IMG + oy) * IMG + ox;
#pragma unroll
    for (int kh = 0; kh < 3; kh++)
#pragma unroll
      for (int kw = 0; kw < 3; kw++) acc[kh * 3 + kw] = fmaf(__ldg(xi + kh * IMG + kw), g, acc[kh * 3 + kw]);
    acc[9] += g;
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
  float acc = 0.f;
  for (long long r = (long long)blockIdx.x * 4 + g; r < rows; r += (long long)gridDim.x * 4) acc += m[r * 64 + c];
  red[g][c] = acc;
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(out + c, red[0][c] + red[1][c] + red[2][c] + red[3][c]);
}

}  // namespace

extern "C" int hvd_conv1_fwd(const float* x, const float* w, const float* bias, float* a1, int B, void* stream) {
  const long long total = (long long)B * mnist::H1 * mnist::H1 * 4;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  conv1_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, w, bias, a1, B);
  return (int)cudaGetLastError();
}

extern "C" int hvd_head(float* h_pre, const float* b1, const float* w2, const float* b2, const long long* labels,
                        float* dh_pre, float* g_w2, float* g_b2, float* g_b1, float* probs, void* step_state, int B,
                        int training, int backward, float keep, void* stream) {
  if (B > 256 || (B * mnist::HID) % 4) return -1;
  HeadParams p{B, training, backward, keep, h_pre, b1, w2, b2, labels, dh_pre, g_w2, g_b2, g_b1, probs,
               reinterpret_cast<mnist::StepState*>(step_state)};
  const int smem = (2 * B * mnist::HID + mnist::HID * mnist::NCLS + B * mnist::NCLS + 64) * 4;
  static int attr_smem = 0;
  if (smem > attr_smem) {
    if (cudaFuncSetAttribute(head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -10;
    attr_smem = smem;
  }
  head_kernel<<<1, 1024, smem, (cudaStream_t)stream>>>(p);
  return (int)cudaGetLastError();
}

extern "C" int hvd_conv1_wgrad(const float* x, const float* da1, float* g_w, float* g_b, int B, void* stream) {
  conv1_wgrad_kernel<<<148 * 4, 256, 0, (cudaStream_t)stream>>>(x, da1, g_w, g_b, B);
  return (int)cudaGetLastError();
}

extern "C" int hvd_colsum64(const float* m, float* out, long long rows, void* stream) {
  colsum64_kernel<<<148 * 2, 256, 0, (cudaStream_t)stream>>>(m, out, rows);
  return (int)cudaGetLastError();
}
