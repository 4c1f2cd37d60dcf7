// Engine state shared between engine.cu (step sequencing) and allreduce.cu (data-parallel hooks).
#pragma once
#include <set>
#include <utility>
#include <vector>

#include "common.cuh"

// which layers run on the tcgen05 kernels (B200EngineConfig.use_tc); a cleared bit selects the fp32 CUDA-core kernel
enum { B200_TC_CONV2_FWD = 1 /* conv1+conv2 forward in one kernel (tc_conv2_fwd.cu) */, B200_TC_FC1_FWD = 2,
       B200_TC_FC1_BWD = 4, B200_TC_CONV2_BWD = 8 /* dgrad + wgrad straight from (dp, argmax): no dense dY */,
       B200_TC_FC1_FP32 = 16 /* fc1 GEMMs as 3xTF32 split-operand products (fp32-class, the reference's SGEMM accuracy) */ };

// Where conv-bucket gradients still sit as per-CTA partial sums when a backward kernel ends.  The kernels
// that consume them (adadelta.cu / allreduce.cu) run one CTA per FIN_CHUNK-float chunk of the bucket and reduce
// the partial rows themselves in a fixed order, so no separate reduction launches sit on the critical
// path.  dst_off: float offset of the segment inside the bucket (a multiple of 256); n: floats per row.
// Those kernels run FIN_COLS float4 columns x FIN_SLICES row slices per CTA: a CTA owns a chunk of FIN_CHUNK = 64
// floats.  256 threads and 4 KB of shared memory are small enough to be co-resident with a persistent conv2
// backward CTA (222 KB of shared memory), so they can run UNDER the conv backward instead of queueing behind it,
// and 292 CTAs for conv2's 18688 floats spread the 11 MB of partial rows over every SM.
constexpr int FIN_COLS = 16, FIN_SLICES = 16, FIN_CHUNK = 4 * FIN_COLS, FIN_THREADS = FIN_COLS * FIN_SLICES;
struct FinalizeSegs {
  struct Seg { const float* src; int nparts, stride, n, dst_off; } seg[2];
};
#ifdef __CUDACC__
// sum over the partial rows of the segment(s) covering floats [off + 4*col, +4) of the bucket; every thread of
// the CTA calls this (col = threadIdx % FIN_COLS, slice = threadIdx / FIN_COLS), the total is returned to slice 0
__device__ __forceinline__ float4 reduce_partial_chunk(const FinalizeSegs& segs, int off, float4 (*red)[FIN_COLS]) {
  const int col = threadIdx.x % FIN_COLS, slice = threadIdx.x / FIN_COLS, nsl = blockDim.x / FIN_COLS;  // <= FIN_SLICES
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const FinalizeSegs::Seg sg = segs.seg[k];
    const int e = off - sg.dst_off + 4 * col;
    if (sg.src && e >= 0 && e < sg.n) {
#pragma unroll 8
      for (int p = slice; p < sg.nparts; p += nsl) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(sg.src + (size_t)p * sg.stride + e));
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    }
  }
  red[slice][col] = acc;
  __syncthreads();
  float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
  if (slice == 0) {
    t = red[0][col];
    for (int q = 1; q < nsl; ++q) { const float4 v = red[q][col]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
  }
  return t;
}
#endif

// adadelta.cu
// conv2's half of the conv bucket on one GPU: partial rows -> gradient buffer (chunks [0, nchunks) of dst)
int b200_reduce_segs(float* dst, int nchunks, const FinalizeSegs& segs, cudaStream_t s);
// single-GPU end of step: chunks [0, first_partial) of the conv bucket take their gradient from g, chunks
// [first_partial, nchunks) reduce `segs` first; then Adadelta on the conv bucket (and on the fc bucket unless
// bucket0_done), and the step counter advances
int b200_adadelta_finalize(float* p, float* g, float* v, float* u, long b1_off, int nchunks, int first_partial,
                           const FinalizeSegs& segs, void* ctrl, float rho, float eps, int advance,
                           unsigned int* done_counter, bool bucket0_done, cudaStream_t s);

struct Engine;
struct EngineHooks {
  // called (inside the step's stream order / graph capture) when gradients are complete:
  //   bucket 0 = fc bucket (in the gradient buffer), 2 = conv2's part of the conv bucket (as partial rows, fin_c2);
  //   the rest of the conv bucket (conv1, fin_c1) is handled by step_end
  int (*bucket_ready)(Engine*, int bucket, cudaStream_t) = nullptr;
  // called once per step after the last bucket: joins side streams, advances the control block
  int (*step_end)(Engine*, int advance, cudaStream_t) = nullptr;
  void (*destroy)(Engine*) = nullptr;
  void* state = nullptr;
};

struct Engine {
  B200EngineConfig cfg{};
  int use_tc = 0, nsm = 148;
  std::vector<void*> owned;
  float *P = nullptr, *G[2] = {nullptr, nullptr}, *V = nullptr, *U = nullptr;
  uint8_t *img[2] = {nullptr, nullptr}, *lbl[2] = {nullptr, nullptr};
  int n[2] = {0, 0};
  int32_t* idx = nullptr;
  bool idx_seq = true;
  int epoch_n = 0;
  uint8_t *stage_img = nullptr, *stage_lbl = nullptr;
  uint8_t *pin_img[2] = {nullptr, nullptr}, *pin_lbl[2] = {nullptr, nullptr};  // pinned host staging (train_step_gather)
  float *a1 = nullptr, *da1 = nullptr, *pooled = nullptr, *dp = nullptr, *h = nullptr, *dh = nullptr,
        *dlogits = nullptr, *loss_i = nullptr, *logp = nullptr, *fc1_part = nullptr, *c1_part = nullptr,
        *c2_part = nullptr, *loss_ring = nullptr, *loss_scalar = nullptr;
  float* c1_gpart = nullptr;      // group-combined partial rows of conv1 wgrad (common.cuh: group_reduce_rows)
  unsigned int* c1_cnt = nullptr;  // their arrival counters
  uint8_t* amax = nullptr;
  int* correct_i = nullptr;
  B200Ctrl *ctrl = nullptr, *ctrl_eval = nullptr;
  unsigned int* done_counter = nullptr;
  double* eval_acc = nullptr;
  void* h_scratch = nullptr;  // 64 B pinned
  float *loss_ring_host = nullptr, *loss_ring_host_dev = nullptr;  // mapped pinned ring + its device alias
  unsigned long long steps = 0;
  bool defer_update = false;  // true: train_step stops after backward (unfused NCCL baseline)
  bool dropout = true;  // false: train without dropout (numerics / DDP-equivalence tests)
  std::set<std::pair<int, int>> warm_batches;  // (batch, mode) pairs that already ran eagerly once
  cudaStream_t copy_stream = nullptr;  // H2D staging for train_step_host (double buffered)
  cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};
  bool consumed_valid[2] = {false, false};
  // Independent kernels of the backward pass run on two branch streams (forked and joined with
  // events, so the whole step is still one capturable graph): br[0] = head wgrad, fc1 wgrad and the
  // fc bucket's Adadelta, br[1] = conv2 wgrad.
  cudaStream_t br[2] = {nullptr, nullptr};
  cudaEvent_t ev_head = nullptr, ev_fcw = nullptr, ev_fcd = nullptr, ev_fcu = nullptr, ev_c2w = nullptr, ev_wg = nullptr;
  bool branches = true;
  EngineHooks hooks;
  FinalizeSegs fin_c2{}, fin_c1{};  // partial rows of conv2 / conv1 wgrad, filled by enqueue_train
  int c2_chunks = 0, conv_chunks = 0;  // FIN_CHUNK-float chunks of conv2.{weight,bias} / of the whole conv bucket
};
