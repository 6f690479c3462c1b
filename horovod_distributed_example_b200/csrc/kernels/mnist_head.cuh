// The classifier head of the MNIST CNN as a device function (reference layers: mnist_keras.py:79-81, loss and metric
// :89-91): bias + ReLU + Dropout(0.5) of Dense(128), Dense(10) + softmax, cross-entropy, accuracy, and in training the
// whole backward of Dense(10) plus the dropout / ReLU backward of Dense(128).  One 256-thread group handles 8 batch
// rows (one warp per row).  Two callers: head_kernel (mnist_small.cu; evaluation, prediction, fallback) and the
// split-K Dense(128) forward GEMM (tc_gemm.cu, EPI_SPLITK_HEAD), whose CTAs run it on the freshly reduced
// pre-activations without a kernel boundary in between.
#pragma once
#include "mnist_common.cuh"

namespace mnist {

struct HeadParams {
  int B, training, backward, rows_per_cta;
  float keep;
  float* h_pre;               // [B,128] split-K sums of dense1 (zeroed again after use)
  const float* b1;            // dense1 bias [128]
  const float* w2;            // dense2 kernel [128,10]
  const float* b2;            // [10]
  const long long* labels;    // [B] int64 class ids
  float* dh_pre;              // [B,128] out: gradient wrt dense1 pre-activation
  float* g_w2; float* g_b2; float* g_b1;   // gradient slots in the arena (accumulated: zeroed by the prologue)
  float* probs;               // [B,10] out (predict / eval), may be null
  StepState* st;
};

constexpr int HEAD_ROWS = 8;  // batch rows per 256-thread group: one warp per row

struct HeadSmem {
  float h[HEAD_ROWS][HID];      // post relu+dropout activations
  float msk[HEAD_ROWS][HID];    // d(h)/d(pre): 0 or 1/keep
  float w2[HID * NCLS];
  float dl[HEAD_ROWS][NCLS + 2];
};

// tid in [0, 256); row_block = which group of 8 rows; sync() = a barrier over exactly these 256 threads.
// Cross-row sums (dW2, db2, db1, loss, accuracy) leave the group as one atomic per output.
template <typename Sync>
__device__ __forceinline__ void head_rows(const HeadParams& p, int row_block, int tid, HeadSmem& sm, Sync sync) {
  const int warp = tid >> 5, lane = tid & 31;
  const int row = row_block * HEAD_ROWS + warp;
  const bool valid = row < p.B;
  const unsigned step = p.st->step, seed = p.st->seed;
  const float inv_keep = 1.0f / p.keep;

  for (int i = tid; i < HID * NCLS; i += 256) sm.w2[i] = __ldg(p.w2 + i);
  {
    // lane owns hidden units 4*lane .. 4*lane+3 of its row
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
      float4* src = reinterpret_cast<float4*>(p.h_pre + (long long)row * HID) + lane;
      v = __ldcg(src);                                         // L2: the sums were made there by red.add
      *src = make_float4(0.f, 0.f, 0.f, 0.f);                  // re-arm the split-K accumulator
    }
    const float4 bb = __ldg(reinterpret_cast<const float4*>(p.b1) + lane);
    uint4 rnd = make_uint4(0, 0, 0, 0);
    if (p.training) rnd = philox4x32_10(make_uint4((unsigned)(row * (HID / 4) + lane), step, 0x1u, 0u), make_uint2(seed, 0x5EEDu));
    const float pre[4] = {v.x + bb.x, v.y + bb.y, v.z + bb.z, v.w + bb.w};
    const unsigned rr[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
    const float scale = p.training ? inv_keep : 1.0f;
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const bool live = valid && (!p.training || u01(rr[e]) < p.keep) && pre[e] > 0.f;
      sm.h[warp][lane * 4 + e] = live ? pre[e] * scale : 0.f;
      sm.msk[warp][lane * 4 + e] = live ? scale : 0.f;
    }
  }
  sync();
  // logits: each lane does its 4 hidden units for all 10 classes, then a warp all-reduce
  float lg[NCLS];
#pragma unroll
  for (int k = 0; k < NCLS; k++) lg[k] = 0.f;
#pragma unroll
  for (int e = 0; e < 4; e++) {
    const float hv = sm.h[warp][lane * 4 + e];
#pragma unroll
    for (int k = 0; k < NCLS; k++) lg[k] = fmaf(hv, sm.w2[(lane * 4 + e) * NCLS + k], lg[k]);
  }
#pragma unroll
  for (int k = 0; k < NCLS; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lg[k] += __shfl_xor_sync(0xffffffffu, lg[k], o);
    lg[k] += __ldg(p.b2 + k);
  }
  // softmax / loss / accuracy (every lane holds the row's 10 logits)
  float mx = lg[0];
  int am = 0;
#pragma unroll
  for (int k = 1; k < NCLS; k++) if (lg[k] > mx) { mx = lg[k]; am = k; }
  float e[NCLS], ssum = 0.f;
#pragma unroll
  for (int k = 0; k < NCLS; k++) { e[k] = expf(lg[k] - mx); ssum += e[k]; }
  const int y = valid ? (int)p.labels[row] : 0;
  const float inv = 1.0f / ssum, invB = 1.0f / (float)p.B;
  float py = 0.f;
#pragma unroll
  for (int k = 0; k < NCLS; k++) {
    const float pk = e[k] * inv;
    if (k == y) py = pk;
    if (lane == k) {
      if (p.probs && valid) p.probs[(long long)row * NCLS + k] = pk;
      sm.dl[warp][k] = valid ? (pk - (k == y ? 1.f : 0.f)) * invB : 0.f;
    }
  }
  if (lane == 0 && valid) {
    const float loss = -logf(fminf(fmaxf(py, 1e-7f), 1.0f - 1e-7f));     // Keras clips probabilities
    atomicAdd(&p.st->last_loss, p.backward ? loss * invB : loss);        // eval mode reports sums
    atomicAdd(&p.st->last_correct, am == y ? (p.backward ? invB : 1.f) : 0.f);   // training: accuracy fraction; eval: count
  }
  sync();
  if (!p.backward) return;
  // dW2[j][k] += sum_rows h[r][j] dl[r][k];  db2[k] += sum_rows dl[r][k]
  for (int o = tid; o < HID * NCLS + NCLS; o += 256) {
    float acc = 0.f;
    if (o < HID * NCLS) {
      const int j = o / NCLS, k = o % NCLS;
#pragma unroll
      for (int r = 0; r < HEAD_ROWS; r++) acc = fmaf(sm.h[r][j], sm.dl[r][k], acc);
      atomicAdd(p.g_w2 + o, acc);
    } else {
      const int k = o - HID * NCLS;
#pragma unroll
      for (int r = 0; r < HEAD_ROWS; r++) acc += sm.dl[r][k];
      atomicAdd(p.g_b2 + k, acc);
    }
  }
  // dh_pre[row][j] = (sum_k dl[row][k] w2[j][k]) * mask;   db1[j] += sum_rows dh_pre[row][j]
  float dh[4];
#pragma unroll
  for (int e4 = 0; e4 < 4; e4++) {
    const int j = lane * 4 + e4;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < NCLS; k++) acc = fmaf(sm.dl[warp][k], sm.w2[j * NCLS + k], acc);
    dh[e4] = acc * sm.msk[warp][j];
  }
  if (valid) reinterpret_cast<float4*>(p.dh_pre + (long long)row * HID)[lane] = make_float4(dh[0], dh[1], dh[2], dh[3]);
  sync();                                  // everyone is done reading h: reuse it for the column sums
#pragma unroll
  for (int e4 = 0; e4 < 4; e4++) sm.h[warp][lane * 4 + e4] = dh[e4];
  sync();
  if (tid < HID) {
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < HEAD_ROWS; r++) acc += sm.h[r][tid];
    atomicAdd(p.g_b1 + tid, acc);
  }
}

}  // namespace mnist
