// Classifier head: everything after the fc1 GEMM, forward and backward, in two kernels.
//
// head_fused   (one warp per sample; four warps share a sample's split-K reduction when there are >= 8 partials)
//     split-K reduce of the fc1 partials + bias + ReLU + dropout(.5)           (mnist_ddp.py:57-59)
//     fc2 (128->10) + log_softmax + NLL                                        (mnist_ddp.py:60-61,71)
//     backward seed: dlogits = (softmax - onehot)/B, dh = (dlogits . W2) * mask
//     eval mode: per-sample NLL and correctness instead of gradients           (mnist_ddp.py:97-99)
// head_wgrad   dW2, db2, db1 and the mean loss: fixed-order reductions over the batch.
//
// N = 10 is below every MMA tile, so this is SIMT by design.  The dropout mask
// is not stored: h = relu(z) * keep * 2, so "h > 0" is exactly "gradient flows" with scale 2.
#include "common.cuh"

using namespace net;

constexpr int DL = 16;  // dlogits row pitch (10 padded to 16 floats)

__global__ void __launch_bounds__(256)
head_fused_kernel(const float* __restrict__ part, int splits, int batch,      // fc1 partials [S][B][128]
                  const float* __restrict__ b1, const float* __restrict__ w2,   // [10][128]
                  const float* __restrict__ b2, const uint8_t* __restrict__ labels,
                  const int32_t* __restrict__ idx, const B200Ctrl* __restrict__ ctrl,
                  float* __restrict__ h,         // [B][128]   post-dropout activation
                  float* __restrict__ dlogits,   // [B][16]    (train)
                  float* __restrict__ dh,        // [B][128]   (train) gradient at fc1's pre-activation
                  float* __restrict__ loss_i,    // [B]        per-sample NLL
                  int* __restrict__ correct_i,   // [B]        (eval) or nullptr
                  float* __restrict__ logp_out,  // [B][10]    optional log-probs, or nullptr
                  int ws)                        // warps per sample (1, 2 or 4)
{
  __shared__ float w2s[NCLS][FC1_OUT];
  __shared__ float b2s[NCLS];
  __shared__ float4 zs[8][32];
  pdl_trigger();
  pdl_wait();
  for (int e = threadIdx.x; e < NCLS * FC1_OUT; e += blockDim.x) w2s[e / FC1_OUT][e % FC1_OUT] = w2[e];
  if (threadIdx.x < NCLS) b2s[threadIdx.x] = b2[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // ws warps share one sample's split-K reduction (each sums every ws-th partial, all its loads in flight at once;
  // the sub-sums are combined in a fixed order through shared memory): with 36 partials a single warp per sample
  // made this kernel a chain of six dependent L2 round trips.
  const int spb = (blockDim.x >> 5) / ws;  // samples per block
  const int sl = warp / ws, sub = warp - sl * ws;
  const int b = blockIdx.x * spb + sl;
  const int train = ctrl->train;

  // ---- fc1 epilogue: reduce split-K partials in a fixed order
  float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  if (b < batch) {
    const float* src = part + (size_t)b * FC1_OUT + lane * 4;
    const size_t pitch = (size_t)batch * FC1_OUT;
    int s = sub;
    for (; s + 8 * ws < splits; s += 9 * ws) {  // nine loads in flight, summed in split order
      float4 p[9];
#pragma unroll
      for (int j = 0; j < 9; ++j) p[j] = *reinterpret_cast<const float4*>(src + (size_t)(s + j * ws) * pitch);
#pragma unroll
      for (int j = 0; j < 9; ++j) { z.x += p[j].x; z.y += p[j].y; z.z += p[j].z; z.w += p[j].w; }
    }
    for (; s < splits; s += ws) {
      const float4 p = *reinterpret_cast<const float4*>(src + (size_t)s * pitch);
      z.x += p.x; z.y += p.y; z.z += p.z; z.w += p.w;
    }
  }
  if (ws > 1) {
    zs[warp][lane] = z;
    __syncthreads();
    if (sub != 0 || b >= batch) return;
    z = zs[warp][lane];
    for (int q = 1; q < ws; ++q) { const float4 v = zs[warp + q][lane]; z.x += v.x; z.y += v.y; z.z += v.z; z.w += v.w; }
  } else if (b >= batch) {
    return;
  }
  {
    const float4 bb = *reinterpret_cast<const float4*>(b1 + lane * 4);
    z.x += bb.x; z.y += bb.y; z.z += bb.z; z.w += bb.w;
  }
  float hv[4] = {fmaxf(z.x, 0.f), fmaxf(z.y, 0.f), fmaxf(z.z, 0.f), fmaxf(z.w, 0.f)};
  if (train & 2) {
    const uint4 r = dropout_block(ctrl->seed, ctrl->step, 2u, (uint32_t)b);  // one block = this sample's 128 units
    const uint32_t m4 = (block_word(r, lane >> 3) >> (4 * (lane & 7))) & 0xfu;  // units 4*lane .. 4*lane+3
    hv[0] = (m4 & 1u) ? hv[0] * 2.f : 0.f;
    hv[1] = (m4 & 2u) ? hv[1] * 2.f : 0.f;
    hv[2] = (m4 & 4u) ? hv[2] * 2.f : 0.f;
    hv[3] = (m4 & 8u) ? hv[3] * 2.f : 0.f;
  }
  *reinterpret_cast<float4*>(h + (size_t)b * FC1_OUT + lane * 4) = make_float4(hv[0], hv[1], hv[2], hv[3]);

  // ---- fc2 + log_softmax
  float logit[NCLS];
#pragma unroll
  for (int c = 0; c < NCLS; ++c) {
    const float4 w = *reinterpret_cast<const float4*>(&w2s[c][lane * 4]);
    float s = hv[0] * w.x;
    s = fmaf(hv[1], w.y, s); s = fmaf(hv[2], w.z, s); s = fmaf(hv[3], w.w, s);
    logit[c] = warp_sum(s) + b2s[c];
  }
  float mx = logit[0];
#pragma unroll
  for (int c = 1; c < NCLS; ++c) mx = fmaxf(mx, logit[c]);
  float se = 0.f;
#pragma unroll
  for (int c = 0; c < NCLS; ++c) se += expf(logit[c] - mx);
  const float lse = mx + logf(se);
  const int sample = idx ? idx[ctrl->pos + b] : ctrl->pos + b;
  const int y = labels[sample];
  float nll = 0.f;
  int arg = 0;
  float best = logit[0];
#pragma unroll
  for (int c = 0; c < NCLS; ++c) {
    if (c == y) nll = lse - logit[c];
    if (logit[c] > best) { best = logit[c]; arg = c; }
  }
  if (lane == 0) {
    loss_i[b] = nll;
    if (correct_i) correct_i[b] = (arg == y);
  }
  if (logp_out && lane < NCLS) {
    float v = 0.f;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) if (c == lane) v = logit[c] - lse;
    logp_out[(size_t)b * NCLS + lane] = v;
  }
  if (!(train & 1)) return;

  // ---- backward seed
  const float invB = 1.0f / (float)batch;
  float dl[NCLS];
#pragma unroll
  for (int c = 0; c < NCLS; ++c) dl[c] = (expf(logit[c] - lse) - (c == y ? 1.f : 0.f)) * invB;
  if (lane < DL) {
    float v = 0.f;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) if (c == lane) v = dl[c];
    dlogits[(size_t)b * DL + lane] = v;
  }
  float g[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int c = 0; c < NCLS; ++c) {
    const float4 w = *reinterpret_cast<const float4*>(&w2s[c][lane * 4]);
    g[0] = fmaf(dl[c], w.x, g[0]); g[1] = fmaf(dl[c], w.y, g[1]);
    g[2] = fmaf(dl[c], w.z, g[2]); g[3] = fmaf(dl[c], w.w, g[3]);
  }
  const float keep_scale = (train & 2) ? 2.f : 1.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) g[j] = hv[j] > 0.f ? g[j] * keep_scale : 0.f;
  *reinterpret_cast<float4*>(dh + (size_t)b * FC1_OUT + lane * 4) = make_float4(g[0], g[1], g[2], g[3]);
}

// grid = 12 blocks x 1024 threads (128 columns x 8 batch slices):
//   blocks 0..9 : dW2[c][:]   block 10 : db1[:]   block 11 : db2[0..9] and the mean loss
// Each thread sums its batch slice serially; slices are combined in a fixed order.
__global__ void __launch_bounds__(1024)
head_wgrad_kernel(const float* __restrict__ h, const float* __restrict__ dlogits, const float* __restrict__ dh,
                  const float* __restrict__ loss_i, int batch, float* __restrict__ dw2, float* __restrict__ db2,
                  float* __restrict__ db1, float* __restrict__ loss_out, const B200Ctrl* __restrict__ ctrl,
                  float* __restrict__ loss_ring, int ring, float* __restrict__ loss_ring_host) {
  __shared__ float red[8][FC1_OUT];
  pdl_trigger();
  pdl_wait();
  const int j = threadIdx.x & 127, sl = threadIdx.x >> 7;
  const int per = (batch + 7) / 8, b0 = sl * per, b1 = min(batch, b0 + per);
  float s = 0.f;
  const int blk = blockIdx.x;
  if (blk < NCLS) {
#pragma unroll 8
    for (int b = b0; b < b1; ++b) s = fmaf(dlogits[(size_t)b * DL + blk], h[(size_t)b * FC1_OUT + j], s);
  } else if (blk == NCLS) {
#pragma unroll 8
    for (int b = b0; b < b1; ++b) s += dh[(size_t)b * FC1_OUT + j];
  } else {
    if (j < NCLS) for (int b = b0; b < b1; ++b) s += dlogits[(size_t)b * DL + j];
    else if (j == NCLS) for (int b = b0; b < b1; ++b) s += loss_i[b];
  }
  red[sl][j] = s;
  __syncthreads();
  if (sl == 0) {
    float t = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += red[q][j];
    if (blk < NCLS) dw2[blk * FC1_OUT + j] = t;
    else if (blk == NCLS) db1[j] = t;
    else if (j < NCLS) db2[j] = t;
    else if (j == NCLS) {
      const float mean = t / (float)batch;
      loss_out[0] = mean;
      if (loss_ring) loss_ring[ctrl->step % (unsigned long long)ring] = mean;
      // zero-copy result read-back: the same value goes straight into mapped, pinned HOST memory
      if (loss_ring_host) loss_ring_host[ctrl->step % (unsigned long long)ring] = mean;
    }
  }
}

// eval accumulation: acc[0] += sum nll (double), acc[1] += #correct, one block, fixed order
__global__ void __launch_bounds__(256)
eval_accumulate_kernel(const float* __restrict__ loss_i, const int* __restrict__ correct_i, int batch,
                       double* __restrict__ acc) {
  __shared__ double sl[256];
  __shared__ int sc[256];
  double l = 0.0;
  int c = 0;
  for (int b = threadIdx.x; b < batch; b += 256) { l += (double)loss_i[b]; c += correct_i[b]; }
  sl[threadIdx.x] = l; sc[threadIdx.x] = c;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { sl[threadIdx.x] += sl[threadIdx.x + o]; sc[threadIdx.x] += sc[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { acc[0] += sl[0]; acc[1] += (double)sc[0]; }
}

extern "C" int b200_head_fused(const float* part, int splits, int batch, const float* b1, const float* w2,
                               const float* b2, const uint8_t* labels, const int32_t* idx, const void* ctrl,
                               float* h, float* dlogits, float* dh, float* loss_i, int* correct_i, float* logp,
                               void* stream) {
  B200_REQUIRE(batch > 0 && splits > 0, "head_fused: bad batch/splits");
  // many partials (small batches): four warps per sample, two samples per CTA; few partials: one warp per sample
  const int ws = splits >= 8 ? 4 : 1, warps = 8, spb = warps / ws;
  B200_CHECK(b200_launch(head_fused_kernel, dim3(ceil_div(batch, spb)), dim3(32 * warps), 0, (cudaStream_t)stream, part, splits,
                         batch, b1, w2, b2, labels, idx, (const B200Ctrl*)ctrl, h, dlogits, dh, loss_i, correct_i, logp, ws));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_head_wgrad(const float* h, const float* dlogits, const float* dh, const float* loss_i,
                               int batch, float* dw2, float* db2, float* db1, float* loss_out, const void* ctrl,
                               float* loss_ring, int ring, float* loss_ring_host, void* stream) {
  B200_REQUIRE(batch > 0, "head_wgrad: bad batch");
  B200_CHECK(b200_launch(head_wgrad_kernel, dim3(NCLS + 2), dim3(1024), 0, (cudaStream_t)stream, h, dlogits, dh, loss_i,
                         batch, dw2, db2, db1, loss_out, (const B200Ctrl*)ctrl, loss_ring, ring, loss_ring_host));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_eval_accumulate(const float* loss_i, const int* correct_i, int batch, double* acc, void* stream) {
  eval_accumulate_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(loss_i, correct_i, batch, acc);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
