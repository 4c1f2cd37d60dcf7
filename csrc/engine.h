// Engine state shared between engine.cu (step sequencing) and allreduce.cu (data-parallel hooks).
#pragma once
#include <set>
#include <utility>
#include <vector>

#include "common.cuh"

enum { B200_TC_CONV2_FWD = 1, B200_TC_FC1_FWD = 2, B200_TC_FC1_BWD = 4, B200_TC_CONV2_BWD = 8, B200_TC_CONV2_V2 = 16, B200_TC_CONV2_V3 = 32,
       B200_TC_CONV2_V4 = 64 /* draft dY-free dgrad (tc_conv2_v4.cu), opt-in only */,
       B200_TC_CONV12_V4 = 128 /* draft fused conv1+conv2 forward (tc_conv2_v4.cu), opt-in only */ };

// Where the conv bucket's gradients still sit as per-CTA partial sums at the end of backward.  The
// last kernel of the step (allreduce.cu: finalize_bucket_kernel) reduces them itself, so no separate
// reduction kernels sit on the critical path.  dst_off: float offset inside the bucket (x256).
constexpr int FIN_SLICES = 16;  // the finalize kernels run 64 float4 columns x FIN_SLICES row slices per CTA
struct FinalizeSegs {
  struct Seg { const float* src; int nparts, stride, n, dst_off; } seg[4];
};

// adadelta.cu: single-GPU end of step = partial reduction of the conv bucket + Adadelta on everything
int b200_adadelta_finalize(float* p, float* g, float* v, float* u, long b1_off, int nchunks, const FinalizeSegs& segs,
                           void* ctrl, float rho, float eps, int advance, unsigned int* done_counter, bool bucket0_done,
                           cudaStream_t s);

struct Engine;
struct EngineHooks {
  // called (inside the step's stream order / graph capture) when a gradient bucket is complete
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
  float *a1 = nullptr, *da1 = nullptr, *pooled = nullptr, *dp = nullptr, *h = nullptr, *dh = nullptr,
        *dlogits = nullptr, *loss_i = nullptr, *logp = nullptr, *fc1_part = nullptr, *c1_part = nullptr,
        *c2_part = nullptr, *dy = nullptr, *db_part = nullptr, *loss_ring = nullptr, *loss_scalar = nullptr;
  float *c1_gpart = nullptr, *db_gpart = nullptr;      // group-combined partial rows (common.cuh: group_reduce_rows)
  unsigned int *c1_cnt = nullptr, *db_cnt = nullptr;   // their arrival counters
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
  cudaEvent_t ev_head = nullptr, ev_fcw = nullptr, ev_fcd = nullptr, ev_fcu = nullptr, ev_dy = nullptr, ev_c2w = nullptr;
  bool branches = true;
  EngineHooks hooks;
  FinalizeSegs fin{};       // filled by enqueue_train when fin_valid
  bool fin_valid = false;
};
