// The training engine: owns device buffers, sequences the kernels of one optimizer step, and
// replays that sequence from a CUDA graph.
//
// It plays the role of the reference's train()/test() loops (mnist_ddp.py:65-105) plus the
// machinery PyTorch runs underneath them — autograd scheduling, the DataLoader/H2D path, the
// DDP reducer hooks, the foreach optimizer — for this one static model.  The step is fully
// device-driven: sample cursor, step counter (dropout offset), lr and the loss ring live in a
// device control block, so the host only enqueues "one more step" and never synchronises.
//
// Kernels of a step (12 launches).  Main stream = the critical path:
//   conv1+conv2 forward in ONE kernel (gather + normalise + conv1 + ReLU in the producer warps, conv2 on tcgen05,
//   bias + ReLU + 2x2 max-pool + arg-max + dropout in the epilogue)
//   -> fc1 GEMM (split-K) -> head_fused (partial sum + bias + ReLU + dropout, fc2, log_softmax, NLL, backward seed)
//   -> fc1 dgrad (+dropout/ReLU mask) -> conv2 wgrad -> conv2 dgrad (both expand (pooled gradient, arg-max) on chip:
//   the dense dY never exists) -> conv1 wgrad
//   -> last kernel: conv-bucket Adadelta (+ conv1's partial rows) + step counter advance.
// Branch 0 (forked after head_fused): head_wgrad -> fc1 wgrad -> [bucket 0 ready] fc-bucket Adadelta
//   (waits for fc1 dgrad, which still reads the old fc1.weight).
// Branch 1 (forked after conv2 wgrad): conv2's per-CTA partial rows -> gradient (-> cross-GPU all-reduce).
// Both branches are joined before the last kernel; forks and joins are events, so the whole step is
// still ONE capturable graph.  With world > 1 the "bucket ready" points launch allreduce.cu's kernels instead:
// fused_allreduce_kernel (+Adadelta) on branch 0, conv2_reduce_kernel on branch 1, conv_finalize_kernel as the last
// kernel.  Gradients are written straight into the flat (symmetric) bucket: there is no bucket copy-in/copy-out
//  and no zero_grad (every wgrad kernel overwrites).
#include <cstdlib>
#include <cstring>
#include <map>
#include <utility>
#include <vector>

#include "common.cuh"
#include "engine.h"

using namespace net;

namespace {

int fc1_splits(int batch) {
  const int tiles = ceil_div(batch, 64) * 2;
  int s = (2 * 148 + tiles / 2) / tiles;
  if (s < 1) s = 1;
  if (s > 36) s = 36;
  while (9216 % (s * 16) != 0) --s;  // equal, 16-aligned K chunks
  return s;
}

template <class T>
int dev_alloc(Engine* E, T** p, size_t count, bool zero = true) {
  void* q = nullptr;
  B200_CHECK(cudaMalloc(&q, count * sizeof(T)));
  if (zero) B200_CHECK(cudaMemset(q, 0, count * sizeof(T)));
  E->owned.push_back(q);
  *p = (T*)q;
  return B200_OK;
}

__global__ void ctrl_rewind_kernel(B200Ctrl* c) { c->pos = 0; }
__global__ void ctrl_lr_kernel(B200Ctrl* c, float lr) { c->lr = lr; }
__global__ void ctrl_train_kernel(B200Ctrl* c, int train) { c->train = train; }
__global__ void zero_eval_kernel(double* acc) { acc[0] = 0.0; acc[1] = 0.0; }

#define TRY(x)                    \
  do {                            \
    int _r = (x);                 \
    if (_r != B200_OK) return _r; \
  } while (0)

// Timeline tracing (B200_TRACE=<step>): in eager mode a timing event is recorded on the launching stream
// after every kernel of that step; the step is held back by a delay kernel until everything is enqueued,
// so the printed times show which kernels overlap on the GPU rather than host launch latency.
struct TraceMark { const char* what; cudaEvent_t ev; int lane; };
std::vector<TraceMark> g_trace;
Engine* g_trace_engine = nullptr;
cudaStream_t g_trace_main = nullptr;

__global__ void trace_delay_kernel(unsigned long long ns) {
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < ns);
}

void trace_mark(Engine* E, const char* what, cudaStream_t st) {
  if (g_trace_engine != E) return;
  cudaEvent_t ev;
  if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cap);
  if (cap == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(ev, st, cudaEventRecordExternal);  // a graph node
  else cudaEventRecord(ev, st);
  g_trace.push_back({what, ev, st == g_trace_main ? 0 : (st == E->br[0] ? 1 : (st == E->br[1] ? 2 : 3))});
}

 void trace_print(long step, int batch, const char* how) {
  fprintf(stderr, "[b200 trace] step %ld, batch %d (%s): end-of-kernel times in us (lane 0 = main stream, 1/2 = branches)\n",
          step, batch, how);
  for (size_t i = 1; i < g_trace.size(); ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, g_trace[0].ev, g_trace[i].ev);
    char name[48];
    snprintf(name, sizeof(name), "%.47s", g_trace[i].what);
    if (char* q = strchr(name, '(')) *q = 0;
    fprintf(stderr, "[b200 trace]   lane %d  %8.2f  %s\n", g_trace[i].lane, ms * 1000.f, name);
  }
}

#define TRYM(x, st)               \
  do {                            \
    int _r = (x);                 \
    if (_r != B200_OK) return _r; \
    trace_mark(E, #x, st);        \
  } while (0)

// Forward through the head.  `ctrl` decides train/eval behaviour of the dropout epilogues.
int enqueue_forward(Engine* E, const uint8_t* images, const uint8_t* labels, const int32_t* idx, const B200Ctrl* ctrl,
                    int batch, bool train, cudaStream_t s) {
  const B200EngineConfig& c = E->cfg;
  float* P = E->P;
  if (E->use_tc & B200_TC_CONV2_FWD) {
    // conv1 is computed inside conv2's producer warps: one kernel from uint8 images to the pooled activation
    TRYM(b200_conv12_fwd_tc(images, idx, ctrl, P + c.offsets[0], P + c.offsets[1], E->a1, P + c.offsets[2],
                             P + c.offsets[3], E->pooled, E->amax, batch, s), s);
  } else {
    TRYM(b200_conv1_fwd(images, idx, ctrl, P + c.offsets[0], P + c.offsets[1], E->a1, batch, s), s);
    TRYM(b200_conv2_fwd_simt(E->a1, P + c.offsets[2], P + c.offsets[3], ctrl, E->pooled, E->amax, batch, s), s);
  }
  const int splits = fc1_splits(batch);
  const int fp32 = (E->use_tc & B200_TC_FC1_FP32) ? 1 : 0;
  if (E->use_tc & B200_TC_FC1_FWD)
    TRYM(b200_tc_gemm_ex(E->pooled, P + c.offsets[4], E->fc1_part, batch, FC1_OUT, FC1_IN, 0, 0, splits, 0, nullptr, 0.f, fp32, s), s);
  else
    TRYM(b200_sgemm(E->pooled, FC1_IN, 1, P + c.offsets[4], 1, FC1_IN, E->fc1_part, batch, FC1_OUT, FC1_IN, splits, 0,
                   nullptr, 0.f, s), s);
  TRYM(b200_head_fused(E->fc1_part, splits, batch, P + c.offsets[5], P + c.offsets[6], P + c.offsets[7], labels, idx,
                      ctrl, E->h, train ? E->dlogits : nullptr, train ? E->dh : nullptr, E->loss_i,
                      train ? nullptr : E->correct_i, train ? nullptr : E->logp, s), s);
  return B200_OK;
}

// One optimizer step.  Stream structure (events = graph edges when captured):
//   s  (critical path): forward -> fc1 dgrad -> conv2 dgrad -> conv1 wgrad -> [join] -> end-of-step kernel
//   br0 (after head)  : head wgrad -> fc1 wgrad -> [after fc1 dgrad] fc-bucket update
//   br1 (after fc1 dgrad): conv2 wgrad -> conv2 half of the conv bucket (partial rows -> gradient, and with data
//                       parallelism the cross-GPU all-reduce of it) under conv2 dgrad / conv1 wgrad
// The end-of-step kernel is left with conv1's 1.3 KB of gradient, the conv bucket's Adadelta and the step counter.
int enqueue_train(Engine* E, const uint8_t* images, const uint8_t* labels, const int32_t* idx, int batch,
                  int advance, cudaStream_t s) {
  const B200EngineConfig& c = E->cfg;
  float* P = E->P;
  float* G = E->G[0];
  const float drop1_scale = E->dropout ? 1.0f / (1.0f - P_DROP1) : 1.0f;
  const int fp32 = (E->use_tc & B200_TC_FC1_FP32) ? 1 : 0;
  TRY(enqueue_forward(E, images, labels, idx, E->ctrl, batch, true, s));
  // Fork: the head's weight gradients and the fc1 wgrad GEMM only feed the optimizer, so they leave the
  // critical path and run next to it.
  const bool par = E->branches && E->br[0] != nullptr;
  cudaStream_t sw = par ? E->br[0] : s;
  if (par) {
    B200_CHECK(cudaEventRecord(E->ev_head, s));
    B200_CHECK(cudaStreamWaitEvent(sw, E->ev_head, 0));
  }
  // fc1: dp = (dh . W1) * (pooled > 0 ? 1/(1-p1) : 0)   -- critical path, enqueued first
  if (E->use_tc & B200_TC_FC1_BWD)
    TRYM(b200_tc_gemm_ex(E->dh, P + c.offsets[4], E->dp, batch, FC1_IN, FC1_OUT, 0, 1, 1, 1, E->pooled, drop1_scale, fp32, s), s);
  else
    TRYM(b200_sgemm(E->dh, FC1_OUT, 1, P + c.offsets[4], FC1_IN, 1, E->dp, batch, FC1_IN, FC1_OUT, 1, 1, E->pooled,
                   drop1_scale, s), s);
  // dW2, db2, db1, mean loss; fc1: dW1[128][9216] = dh^T . pooled
  TRYM(b200_head_wgrad(E->h, E->dlogits, E->dh, E->loss_i, batch, G + c.offsets[6], G + c.offsets[7], G + c.offsets[5],
                      E->loss_scalar, E->ctrl, E->loss_ring, B200_LOSS_RING, E->loss_ring_host_dev, sw), sw);
  if (E->use_tc & B200_TC_FC1_BWD)
    TRYM(b200_tc_gemm_ex(E->dh, E->pooled, G + c.offsets[4], FC1_OUT, FC1_IN, batch, 1, 1, 1, 0, nullptr, 0.f, fp32, sw), sw);
  else
    TRYM(b200_sgemm(E->dh, 1, FC1_OUT, E->pooled, FC1_IN, 1, G + c.offsets[4], FC1_OUT, FC1_IN, batch, 1, 0, nullptr,
                   0.f, sw), sw);
  const bool ddp = E->hooks.bucket_ready != nullptr;
  const bool own_update = !ddp && !E->defer_update;
  bool fc_updated = false, fc_pending = false;
  if (par) {
    B200_CHECK(cudaEventRecord(E->ev_fcd, s));  // fc1 dgrad done: dp is ready, the old fc1.weight has been read
    if (E->defer_update) {
      B200_CHECK(cudaEventRecord(E->ev_fcw, sw));  // the caller reduces/updates: the fc bucket must be complete
      fc_pending = true;
    } else {
      // The fc bucket (98 % of the parameters) is complete on the branch.  Its update -- Adadelta on one
      // GPU, the fused all-reduce + Adadelta kernel with data parallelism -- stays on the branch, under
      // the conv backward; it may only start once fc1 dgrad has read the old fc1.weight.
      B200_CHECK(cudaStreamWaitEvent(sw, E->ev_fcd, 0));
      if (own_update) TRYM(b200_adadelta(P, G, E->V, E->U, c.bucket_begin[1], E->ctrl, 0.9f, 1e-6f, -1, nullptr, sw), sw);
      else TRYM(E->hooks.bucket_ready(E, 0, sw), sw);
      B200_CHECK(cudaEventRecord(E->ev_fcu, sw));
      fc_updated = true;
    }
  } else if (ddp) {
    TRYM(E->hooks.bucket_ready(E, 0, s), s);  // bucket 0 (fc1.*, fc2.*) is complete here
    fc_updated = true;
  }
  // ---- conv2 backward; both kernels read (dp, argmax) directly.
  // wgrad and dgrad each need nearly all shared memory of an SM, so they never overlap: they run one after the
  // other in whatever order the block scheduler picks, and with two ranks picking differently every cross-GPU
  // barrier waits for the slower choice (measured).  The order is therefore fixed: wgrad FIRST on the main stream,
  // then dgrad; what follows wgrad (partial rows -> gradient -> cross-GPU all-reduce of conv2's 74 KB) forks onto
  // branch 1 and has all of dgrad + conv1 wgrad (40 us) to hide under.  B200_CONV2_ORDER=0 restores the unordered
  // two-stream form.
  static const bool ordered = !(getenv("B200_CONV2_ORDER") && atoi(getenv("B200_CONV2_ORDER")) == 0);
  const long b1 = c.bucket_begin[1];
  cudaStream_t sc = par ? E->br[1] : s;         // stream of the conv2 reduction
  cudaStream_t swg = (par && !ordered) ? sc : s;  // stream of conv2 wgrad
  if (par && !ordered) B200_CHECK(cudaStreamWaitEvent(sc, E->ev_fcd, 0));
  int np2 = 0;
  for (int pass = 0; pass < 2; ++pass) {
    // unordered form: the critical-path kernel (dgrad) is enqueued first
    const bool do_wgrad = (pass == 0) != (par && !ordered);
    if (E->use_tc & B200_TC_CONV2_BWD) {
      if (do_wgrad) {
        TRYM(b200_conv2_wgrad_tc(E->dp, E->amax, E->a1, E->c2_part, nullptr, nullptr, batch, swg), swg);
        np2 = b200_conv2_wgrad_tc_parts(batch);
      } else {
        TRYM(b200_conv2_dgrad_tc(E->dp, E->amax, P + c.offsets[2], E->a1, E->da1, batch, s), s);
      }
    } else {
      if (do_wgrad) {
        TRYM(b200_conv2_wgrad_simt(E->dp, E->amax, E->a1, E->c2_part, nullptr, nullptr, batch, swg), swg);
        np2 = min(9 * batch, E->nsm);
      } else {
        TRYM(b200_conv2_dgrad_simt(E->dp, E->amax, P + c.offsets[2], E->a1, E->da1, batch, s), s);
      }
    }
    if (do_wgrad && par && ordered) {  // fork: the reduction leaves the main stream right after wgrad
      B200_CHECK(cudaEventRecord(E->ev_wg, s));
      B200_CHECK(cudaStreamWaitEvent(sc, E->ev_wg, 0));
    }
  }
  const int c2row = 9 * C2 * C1 + C2;
  E->fin_c2.seg[0] = {E->c2_part, np2, c2row, 9 * C2 * C1, (int)(c.offsets[2] - b1)};
  E->fin_c2.seg[1] = {E->c2_part + 9 * C2 * C1, np2, c2row, C2, (int)(c.offsets[3] - b1)};
  if (ddp) TRYM(E->hooks.bucket_ready(E, 2, sc), sc);  // partial rows -> gradient -> cross-GPU all-reduce
  else TRYM(b200_reduce_segs(G + b1, E->c2_chunks, E->fin_c2, sc), sc);
  if (par) B200_CHECK(cudaEventRecord(E->ev_c2w, sc));
  // ---- conv1 wgrad (critical path), left as group-combined partial rows for the end-of-step kernel
  const bool fuse = ddp || !E->defer_update;
  if (fuse) {
    TRYM(b200_conv1_wgrad_grouped(images, idx, E->ctrl, E->da1, E->c1_part, E->c1_gpart, E->c1_cnt, batch, s), s);
    const int ng = b200_group_count(2 * batch);
    E->fin_c1.seg[0] = {E->c1_gpart, ng, 320, 288, (int)(c.offsets[0] - b1)};
    E->fin_c1.seg[1] = {E->c1_gpart + 288, ng, 320, 32, (int)(c.offsets[1] - b1)};
  } else {
    TRYM(b200_conv1_wgrad(images, idx, E->ctrl, E->da1, E->c1_part, G + c.offsets[0], G + c.offsets[1], batch, s), s);
  }
  // Join: everything the step's last kernel reads must be complete
  if (par) B200_CHECK(cudaStreamWaitEvent(s, E->ev_c2w, 0));
  if (fc_updated && par) B200_CHECK(cudaStreamWaitEvent(s, E->ev_fcu, 0));
  if (fc_pending) B200_CHECK(cudaStreamWaitEvent(s, E->ev_fcw, 0));
  if (ddp) {
    TRYM(E->hooks.step_end(E, advance, s), s);
  } else if (!E->defer_update) {
    // bucket 0 occupies floats [0, bucket_begin[1]) of the flat layout, bucket 1 the rest
    TRYM(b200_adadelta_finalize(P, G, E->V, E->U, b1, E->conv_chunks, E->c2_chunks, E->fin_c1, E->ctrl, 0.9f, 1e-6f,
                               advance, E->done_counter, fc_updated, s), s);
  }
  return B200_OK;
}

struct GraphKey {
  int batch, mode;
  bool operator<(const GraphKey& o) const { return batch != o.batch ? batch < o.batch : mode < o.mode; }
};
struct GraphVal {
  cudaGraphExec_t exec;
  int kernels;
};
std::map<std::pair<Engine*, GraphKey>, GraphVal> g_graphs;

int run_step(Engine* E, const uint8_t* images, const uint8_t* labels, const int32_t* idx, int batch, int advance,
             int mode, int use_graph, cudaStream_t s) {
  if (!use_graph) {
    static const long trace_step = getenv("B200_TRACE") ? atol(getenv("B200_TRACE")) : -1;
    if (trace_step < 0 || (long)E->steps != trace_step) return enqueue_train(E, images, labels, idx, batch, advance, s);
    B200_CHECK(cudaDeviceSynchronize());
    g_trace.clear();
    g_trace_engine = E;
    g_trace_main = s;
    trace_delay_kernel<<<1, 1, 0, s>>>(300000ull);
    trace_mark(E, "start", s);
    int r = enqueue_train(E, images, labels, idx, batch, advance, s);
    g_trace_engine = nullptr;
    B200_CHECK(cudaDeviceSynchronize());
    trace_print(trace_step, batch, "eager");
    for (auto& m : g_trace) cudaEventDestroy(m.ev);
    g_trace.clear();
    return r;
  }
  static const long trace_step_g = getenv("B200_TRACE") ? atol(getenv("B200_TRACE")) : -1;
  mode += (E->dropout ? 0 : 8) + (E->defer_update ? 16 : 0);
  auto key = std::make_pair(E, GraphKey{batch, mode});
  auto it = g_graphs.find(key);
  if (it == g_graphs.end()) {
    // First use: run once eagerly (lazy attribute set-up, tensor-map creation, error surfacing),
    // capture on the next call.
    if (E->warm_batches.insert({batch, mode}).second)
      return enqueue_train(E, images, labels, idx, batch, advance, s);
    cudaGraph_t graph = nullptr;
    const long long before = b200_launch_count();
    B200_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    const bool traced = trace_step_g >= 0 && g_trace.empty();  // timing events become nodes of the first captured graph
    if (traced) {
      g_trace_engine = E;
      g_trace_main = s;
      trace_mark(E, "start", s);
    }
    int r = enqueue_train(E, images, labels, idx, batch, advance, s);
    g_trace_engine = nullptr;
    cudaError_t ce = cudaStreamEndCapture(s, &graph);
    if (r != B200_OK) {
      if (graph) cudaGraphDestroy(graph);
      return r;
    }
    B200_CHECK(ce);
    GraphVal v{nullptr, (int)(b200_launch_count() - before)};
    b200_count_launch(-v.kernels);  // capture did not execute anything
    B200_CHECK(cudaGraphInstantiate(&v.exec, graph, 0));
    cudaGraphDestroy(graph);
    it = g_graphs.emplace(key, v).first;
  }
  B200_CHECK(cudaGraphLaunch(it->second.exec, s));
  b200_count_launch(it->second.kernels);
  if (trace_step_g >= 0 && (long)E->steps == trace_step_g && !g_trace.empty()) {
    B200_CHECK(cudaStreamSynchronize(s));
    trace_print(trace_step_g, batch, "graph replay");
  }
  return B200_OK;
}

}  // namespace

extern "C" int b200_engine_create(const B200EngineConfig* cfg, void** out) {
  B200_REQUIRE(cfg && out, "engine_create: null argument");
  B200_REQUIRE(cfg->max_batch > 0 && cfg->param_floats > 0 && cfg->param_floats % 4 == 0,
               "engine_create: bad max_batch/param_floats");
  B200_CHECK(cudaSetDevice(cfg->device));
  Engine* E = new Engine();
  E->cfg = *cfg;
  E->use_tc = cfg->use_tc;
  E->nsm = b200_sm_count();
  // the conv bucket is [conv2.weight | conv2.bias | conv1.weight | conv1.bias] in 256-float chunks (ParamLayout)
  B200_REQUIRE(cfg->offsets[2] == cfg->bucket_begin[1] && cfg->offsets[3] > cfg->offsets[2] &&
               cfg->offsets[0] > cfg->offsets[3] && cfg->offsets[1] > cfg->offsets[0] &&
               (cfg->bucket_end[1] - cfg->bucket_begin[1]) % 256 == 0 && (cfg->offsets[0] - cfg->bucket_begin[1]) % 256 == 0,
               "engine_create: unexpected conv bucket layout");
  E->c2_chunks = (int)((cfg->offsets[0] - cfg->bucket_begin[1]) / FIN_CHUNK);
  E->conv_chunks = (int)((cfg->bucket_end[1] - cfg->bucket_begin[1]) / FIN_CHUNK);
  const size_t B = (size_t)cfg->max_batch, T = (size_t)cfg->param_floats;
  int r = B200_OK;
#define A(ptr, count) if (r == B200_OK) r = dev_alloc(E, &(ptr), (count))
  if (cfg->params) E->P = cfg->params; else A(E->P, T);
  for (int i = 0; i < 2; ++i) {
    if (cfg->grads[i]) E->G[i] = cfg->grads[i];
    else if (i == 0 || cfg->world > 1) { A(E->G[i], T); }
    else E->G[i] = E->G[0];
  }
  if (cfg->sq_avg) E->V = cfg->sq_avg; else A(E->V, T);
  if (cfg->acc_delta) E->U = cfg->acc_delta; else A(E->U, T);
  for (int sp = 0; sp < 2; ++sp) {
    const size_t n = sp == 0 ? cfg->train_n : cfg->test_n;
    E->n[sp] = 0;
    A(E->img[sp], n * 784 + 16);
    A(E->lbl[sp], n + 16);
  }
  A(E->idx, (size_t)cfg->train_n + 16);
  A(E->stage_img, 2 * B * 784);
  A(E->stage_lbl, 2 * B);
  A(E->a1, B * H1 * W1 * C1);
  A(E->da1, B * H1 * W1 * C1);
  A(E->pooled, B * FC1_IN);
  A(E->dp, B * FC1_IN);
  A(E->amax, B * FC1_IN);
  A(E->h, B * FC1_OUT);
  A(E->dh, B * FC1_OUT);
  A(E->dlogits, B * 16);
  A(E->loss_i, B);
  A(E->correct_i, B);
  A(E->logp, B * NCLS);
  size_t part_cap = 0;
  for (int b = 1; b <= cfg->max_batch; ++b) part_cap = max(part_cap, (size_t)fc1_splits(b) * b * FC1_OUT);
  A(E->fc1_part, part_cap);
  A(E->c1_part, 2 * B * 320);
  A(E->c1_gpart, (size_t)b200_group_count(2 * cfg->max_batch) * 320);
  A(E->c1_cnt, (size_t)b200_group_count(2 * cfg->max_batch));
  A(E->c2_part, (size_t)E->nsm * (9 * C2 * C1 + C2));
  A(E->loss_ring, (size_t)B200_LOSS_RING);
  A(E->loss_scalar, 4);
  A(E->ctrl, 1);
  A(E->ctrl_eval, 1);
  A(E->done_counter, 4);
  A(E->eval_acc, 2);
#undef A
  if (r != B200_OK) { b200_engine_destroy(E); return r; }
  B200_CHECK(cudaMallocHost((void**)&E->h_scratch, 64));
  B200_CHECK(cudaHostAlloc((void**)&E->loss_ring_host, B200_LOSS_RING * sizeof(float), cudaHostAllocMapped));
  memset(E->loss_ring_host, 0, B200_LOSS_RING * sizeof(float));
  B200_CHECK(cudaHostGetDevicePointer((void**)&E->loss_ring_host_dev, E->loss_ring_host, 0));
  E->branches = !(getenv("B200_NO_BRANCH") && atoi(getenv("B200_NO_BRANCH")) == 1);
  // Branch streams get the LOWEST priority and the engine's own stream (created by the caller) should have a higher
  // one: where a branch kernel and a critical-path kernel compete for SMs (conv2 wgrad vs conv2 dgrad: both need
  // nearly all shared memory of an SM, so they run one after the other), the block scheduler starts the
  // critical-path kernel first.  Stream priorities are recorded in the captured graph's kernel nodes.
  int prio_lo = 0, prio_hi = 0;
  B200_CHECK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  static const bool flat = getenv("B200_FLAT_PRIORITY") != nullptr;  // diagnostics
  for (int i = 0; i < 2; ++i)
    B200_CHECK(cudaStreamCreateWithPriority(&E->br[i], cudaStreamNonBlocking, flat ? prio_hi : prio_lo));
  for (cudaEvent_t* e : {&E->ev_head, &E->ev_fcw, &E->ev_fcd, &E->ev_fcu, &E->ev_c2w, &E->ev_wg})
    B200_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  const float inv_w = 1.0f / (float)(cfg->world > 0 ? cfg->world : 1);
  TRY(b200_ctrl_set(E->ctrl, 0, cfg->seed, 0, 3, 1.0f, inv_w, nullptr));
  TRY(b200_ctrl_set(E->ctrl_eval, 0, cfg->seed, 0, 0, 0.0f, inv_w, nullptr));
  B200_CHECK(cudaDeviceSynchronize());
  *out = E;
  return B200_OK;
}

extern "C" int b200_engine_destroy(void* engine) {
  Engine* E = (Engine*)engine;
  if (!E) return B200_OK;
  cudaDeviceSynchronize();
  for (auto it = g_graphs.begin(); it != g_graphs.end();) {
    if (it->first.first == E) { cudaGraphExecDestroy(it->second.exec); it = g_graphs.erase(it); }
    else ++it;
  }
  if (E->hooks.destroy) E->hooks.destroy(E);
  for (int i = 0; i < 2; ++i) if (E->br[i]) cudaStreamDestroy(E->br[i]);
  for (cudaEvent_t e : {E->ev_head, E->ev_fcw, E->ev_fcd, E->ev_fcu, E->ev_c2w, E->ev_wg}) if (e) cudaEventDestroy(e);
  if (E->copy_stream) {
    cudaStreamDestroy(E->copy_stream);
    for (int i = 0; i < 2; ++i) { cudaEventDestroy(E->ev_copied[i]); cudaEventDestroy(E->ev_consumed[i]); }
  }
  for (void* p : E->owned) cudaFree(p);
  for (int i = 0; i < 2; ++i) {
    if (E->pin_img[i]) cudaFreeHost(E->pin_img[i]);
    if (E->pin_lbl[i]) cudaFreeHost(E->pin_lbl[i]);
  }
  if (E->h_scratch) cudaFreeHost(E->h_scratch);
  if (E->loss_ring_host) cudaFreeHost(E->loss_ring_host);
  delete E;
  return B200_OK;
}

extern "C" int b200_engine_buffer(void* engine, int which, void** ptr, size_t* bytes) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && ptr && bytes, "engine_buffer: null argument");
  const size_t B = (size_t)E->cfg.max_batch, T = (size_t)E->cfg.param_floats;
  switch (which) {
    case B200_BUF_PARAMS: *ptr = E->P; *bytes = T * 4; break;
    case B200_BUF_GRADS0: *ptr = E->G[0]; *bytes = T * 4; break;
    case B200_BUF_GRADS1: *ptr = E->G[1]; *bytes = T * 4; break;
    case B200_BUF_SQ_AVG: *ptr = E->V; *bytes = T * 4; break;
    case B200_BUF_ACC_DELTA: *ptr = E->U; *bytes = T * 4; break;
    case B200_BUF_A1: *ptr = E->a1; *bytes = B * H1 * W1 * C1 * 4; break;
    case B200_BUF_POOLED: *ptr = E->pooled; *bytes = B * FC1_IN * 4; break;
    case B200_BUF_AMAX: *ptr = E->amax; *bytes = B * FC1_IN; break;
    case B200_BUF_H: *ptr = E->h; *bytes = B * FC1_OUT * 4; break;
    case B200_BUF_DLOGITS: *ptr = E->dlogits; *bytes = B * 16 * 4; break;
    case B200_BUF_DH: *ptr = E->dh; *bytes = B * FC1_OUT * 4; break;
    case B200_BUF_DP: *ptr = E->dp; *bytes = B * FC1_IN * 4; break;
    case B200_BUF_DA1: *ptr = E->da1; *bytes = B * H1 * W1 * C1 * 4; break;
    case B200_BUF_LOSS_I: *ptr = E->loss_i; *bytes = B * 4; break;
    case B200_BUF_LOSS_RING: *ptr = E->loss_ring; *bytes = B200_LOSS_RING * 4; break;
    case B200_BUF_CTRL: *ptr = E->ctrl; *bytes = sizeof(B200Ctrl); break;
    case B200_BUF_LOGP: *ptr = E->logp; *bytes = B * NCLS * 4; break;
    case B200_BUF_STAGE_IMG: *ptr = E->stage_img; *bytes = 2 * B * 784; break;
    case B200_BUF_STAGE_LBL: *ptr = E->stage_lbl; *bytes = 2 * B; break;
    case B200_BUF_FC1_PART: *ptr = E->fc1_part; *bytes = 0; break;
    default: B200_REQUIRE(false, "engine_buffer: unknown buffer id %d", which);
  }
  return B200_OK;
}

extern "C" int b200_engine_load_data(void* engine, int split, const uint8_t* images_host, const uint8_t* labels_host,
                                     int n) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && (split == 0 || split == 1), "engine_load_data: bad split");
  B200_REQUIRE(n > 0 && n <= (split == 0 ? E->cfg.train_n : E->cfg.test_n), "engine_load_data: n=%d exceeds capacity", n);
  B200_CHECK(cudaMemcpy(E->img[split], images_host, (size_t)n * 784, cudaMemcpyHostToDevice));
  B200_CHECK(cudaMemcpy(E->lbl[split], labels_host, (size_t)n, cudaMemcpyHostToDevice));
  E->n[split] = n;
  return B200_OK;
}

extern "C" int b200_engine_begin_epoch(void* engine, const int32_t* idx_host, int n, void* stream) {
  Engine* E = (Engine*)engine;
  cudaStream_t s = (cudaStream_t)stream;
  B200_REQUIRE(E && n > 0 && n <= E->cfg.train_n, "engine_begin_epoch: n=%d out of range", n);
  if (idx_host) {
    B200_CHECK(cudaMemcpyAsync(E->idx, idx_host, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    E->idx_seq = false;
  } else {
    E->idx_seq = true;
  }
  E->epoch_n = n;
  ctrl_rewind_kernel<<<1, 1, 0, s>>>(E->ctrl);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_engine_set_lr(void* engine, float lr, void* stream) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E, "engine_set_lr: null engine");
  ctrl_lr_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(E->ctrl, lr);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_engine_set_dropout(void* engine, int on, void* stream) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E, "engine_set_dropout: null engine");
  E->dropout = on != 0;
  ctrl_train_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(E->ctrl, on ? 3 : 1);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Baseline ("--comm nccl") support: with defer = 1 a train step stops after backward; the caller
// reduces the gradient buffer itself (e.g. ncclAllReduce) and then calls b200_engine_apply_update.
extern "C" int b200_engine_set_defer_update(void* engine, int defer) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && !E->hooks.bucket_ready, "engine_set_defer_update: not available with fused data-parallel hooks");
  E->defer_update = defer != 0;
  return B200_OK;
}

extern "C" int b200_engine_apply_update(void* engine, int advance, void* stream) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E, "engine_apply_update: null engine");
  return b200_adadelta(E->P, E->G[0], E->V, E->U, E->cfg.param_floats, E->ctrl, 0.9f, 1e-6f, advance, E->done_counter,
                       stream);
}

extern "C" int b200_engine_train_step(void* engine, int batch, int use_graph, void* stream) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && batch > 0 && batch <= E->cfg.max_batch, "engine_train_step: batch %d out of range", batch);
  B200_REQUIRE(E->n[0] > 0, "engine_train_step: no training data loaded");
  const int32_t* idx = E->idx_seq ? nullptr : E->idx;
  TRY(run_step(E, E->img[0], E->lbl[0], idx, batch, batch, E->idx_seq ? 1 : 0, use_graph, (cudaStream_t)stream));
  E->steps++;
  return B200_OK;
}

extern "C" int b200_engine_train_step_host(void* engine, const uint8_t* images_host, const uint8_t* labels_host,
                                           int batch, int use_graph, void* stream) {
  Engine* E = (Engine*)engine;
  cudaStream_t s = (cudaStream_t)stream;
  B200_REQUIRE(E && batch > 0 && batch <= E->cfg.max_batch, "engine_train_step_host: batch %d out of range", batch);
  const int par = (int)(E->steps & 1);
  uint8_t* di = E->stage_img + (size_t)par * E->cfg.max_batch * 784;
  uint8_t* dl = E->stage_lbl + (size_t)par * E->cfg.max_batch;
  // The copy runs on its own stream into one of two staging buffers, so the H2D transfer of step
  // i+1 overlaps the compute of step i; events keep a buffer from being overwritten while a step
  // still reads it.
  if (!E->copy_stream) {
    B200_CHECK(cudaStreamCreateWithFlags(&E->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      B200_CHECK(cudaEventCreateWithFlags(&E->ev_copied[i], cudaEventDisableTiming));
      B200_CHECK(cudaEventCreateWithFlags(&E->ev_consumed[i], cudaEventDisableTiming));
    }
  }
  if (E->consumed_valid[par]) B200_CHECK(cudaStreamWaitEvent(E->copy_stream, E->ev_consumed[par], 0));
  B200_CHECK(cudaMemcpyAsync(di, images_host, (size_t)batch * 784, cudaMemcpyHostToDevice, E->copy_stream));
  B200_CHECK(cudaMemcpyAsync(dl, labels_host, (size_t)batch, cudaMemcpyHostToDevice, E->copy_stream));
  B200_CHECK(cudaEventRecord(E->ev_copied[par], E->copy_stream));
  B200_CHECK(cudaStreamWaitEvent(s, E->ev_copied[par], 0));
  // staged batches are addressed from sample 0: the cursor must not move (advance = 0)
  TRY(run_step(E, di, dl, nullptr, batch, 0, 2 + par, use_graph, s));
  B200_CHECK(cudaEventRecord(E->ev_consumed[par], s));
  E->consumed_valid[par] = true;
  E->steps++;
  return B200_OK;
}

// The data-loader path in one call: gather `batch` samples of a HOST-resident uint8 dataset by index into pinned
// staging memory (what the reference's DataLoader worker + collate + pin thread do, mnist_ddp.py:144-168), copy
// them to the device on the copy stream, run the step.  Replaces per-sample PIL/ToTensor/Normalize work with a
// 784-byte memcpy per sample; normalisation happens inside the forward kernel.  The host blocks only if the pinned
// slot of two steps ago has not been copied out yet.
extern "C" int b200_engine_train_step_gather(void* engine, const uint8_t* images_host, const uint8_t* labels_host,
                                             int n_samples, const int32_t* idx_host, int batch, int use_graph,
                                             void* stream) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && images_host && labels_host && idx_host && n_samples > 0 && batch > 0 && batch <= E->cfg.max_batch,
               "engine_train_step_gather: bad arguments (batch %d, max %d)", batch, E ? E->cfg.max_batch : 0);
  for (int b = 0; b < batch; ++b)  // a sampler bug must not become a wild host read
    B200_REQUIRE(idx_host[b] >= 0 && idx_host[b] < n_samples,
                 "engine_train_step_gather: index %d at position %d is outside the dataset of %d samples", idx_host[b], b,
                 n_samples);
  const int par = (int)(E->steps & 1);
  if (!E->pin_img[0]) {
    for (int i = 0; i < 2; ++i) {
      B200_CHECK(cudaHostAlloc((void**)&E->pin_img[i], (size_t)E->cfg.max_batch * 784, cudaHostAllocDefault));
      B200_CHECK(cudaHostAlloc((void**)&E->pin_lbl[i], (size_t)E->cfg.max_batch, cudaHostAllocDefault));
    }
  }
  if (E->copy_stream && E->consumed_valid[par]) B200_CHECK(cudaEventSynchronize(E->ev_copied[par]));
  uint8_t* pi = E->pin_img[par];
  uint8_t* pl = E->pin_lbl[par];
  for (int b = 0; b < batch; ++b) {
    const size_t src = (size_t)idx_host[b];
    memcpy(pi + (size_t)b * 784, images_host + src * 784, 784);
    pl[b] = labels_host[src];
  }
  return b200_engine_train_step_host(engine, pi, pl, batch, use_graph, stream);
}

extern "C" int b200_engine_eval_reset(void* engine, void* stream) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E, "engine_eval_reset: null engine");
  zero_eval_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(E->eval_acc);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_engine_eval_batch(void* engine, int split, int first, int batch, void* stream) {
  Engine* E = (Engine*)engine;
  cudaStream_t s = (cudaStream_t)stream;
  B200_REQUIRE(E && (split == 0 || split == 1), "engine_eval_batch: bad split");
  B200_REQUIRE(batch > 0 && batch <= E->cfg.max_batch && first >= 0 && first + batch <= E->n[split],
               "engine_eval_batch: range [%d, %d) outside split of %d samples (max batch %d)", first, first + batch,
               E->n[split], E->cfg.max_batch);
  TRY(b200_ctrl_set(E->ctrl_eval, 0, E->cfg.seed, first, 0, 0.f, 1.f, s));
  TRY(enqueue_forward(E, E->img[split], E->lbl[split], nullptr, E->ctrl_eval, batch, false, s));
  TRY(b200_eval_accumulate(E->loss_i, E->correct_i, batch, E->eval_acc, s));
  return B200_OK;
}

extern "C" int b200_engine_eval_result(void* engine, double* nll_sum_host, long long* correct_host, void* stream) {
  Engine* E = (Engine*)engine;
  cudaStream_t s = (cudaStream_t)stream;
  B200_REQUIRE(E && nll_sum_host && correct_host, "engine_eval_result: null argument");
  double* hs = (double*)E->h_scratch;
  B200_CHECK(cudaMemcpyAsync(hs, E->eval_acc, 16, cudaMemcpyDeviceToHost, s));
  B200_CHECK(cudaStreamSynchronize(s));
  *nll_sum_host = hs[0];
  *correct_host = (long long)(hs[1] + 0.5);
  return B200_OK;
}

extern "C" int b200_engine_read_loss(void* engine, unsigned long long step, float* loss_host, void* stream) {
  Engine* E = (Engine*)engine;
  cudaStream_t s = (cudaStream_t)stream;
  B200_REQUIRE(E && loss_host, "engine_read_loss: null argument");
  float* hs = (float*)E->h_scratch + 8;
  B200_CHECK(cudaMemcpyAsync(hs, E->loss_ring + (step % B200_LOSS_RING), 4, cudaMemcpyDeviceToHost, s));
  B200_CHECK(cudaStreamSynchronize(s));
  *loss_host = *hs;
  return B200_OK;
}

extern "C" int b200_engine_host_loss_ring(void* engine, float** ring_host) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && ring_host, "engine_host_loss_ring: null argument");
  *ring_host = E->loss_ring_host;
  return B200_OK;
}

extern "C" int b200_engine_step_count(void* engine, unsigned long long* steps_host) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && steps_host, "engine_step_count: null argument");
  *steps_host = E->steps;
  return B200_OK;
}

extern "C" int b200_engine_launches_per_step(void* engine, int batch, int* n_host) {
  Engine* E = (Engine*)engine;
  B200_REQUIRE(E && n_host, "engine_launches_per_step: null argument");
  *n_host = -1;
  for (auto& kv : g_graphs)
    if (kv.first.first == E && kv.first.second.batch == batch) *n_host = kv.second.kernels;
  return B200_OK;
}
