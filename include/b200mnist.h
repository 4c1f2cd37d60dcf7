/* libb200mnist — C ABI of the B200-native MNIST-CNN training runtime.
 *
 * Every function returns B200_OK (0) or a negative error code; b200_last_error() gives the
 * message of the last failure on the calling thread.  Pointers are DEVICE pointers unless the
 * parameter name ends in _host.  `stream` is a cudaStream_t passed as void*.  Nothing here
 * synchronises the device unless its comment says so.  All kernels are compiled for sm_100a only.
 *
 * Citations (file:line) refer to the reference repository FlyingAnt2018/pytorch_mnist_ddp; the
 * reference itself has no native code — each entry point replaces the library kernels that the
 * cited Python call site reaches.
 */
#ifndef B200MNIST_H_
#define B200MNIST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { B200_OK = 0, B200_ERR_ARG = -1, B200_ERR_CUDA = -2, B200_ERR_STATE = -3, B200_ERR_UNSUPPORTED = -4 };

/* ---- runtime ------------------------------------------------------------------------------ */
const char* b200_last_error(void);
int b200_version(void);
/* 1 when the library was built with the tcgen05 kernels (always, for sm_100a) */
int b200_has_tc(void);
/* number of kernels this library has launched in this process (graph replays included) */
long long b200_launch_count(void);
/* write the device control block (struct B200Ctrl in csrc/common.cuh) asynchronously */
int b200_ctrl_set(void* ctrl, unsigned long long step, unsigned long long seed, int pos, int train, float lr,
                  float inv_world, void* stream);
int b200_ctrl_bytes(void);

/* ---- data + conv1 (mnist_ddp.py:50-51,153-160; SURVEY N1,N3,N14,N21) --------------------------- */
/* a1[b] = relu(conv1(normalize(images[idx ? idx[pos+b] : pos+b]))), NHWC [B][26][26][32] */
int b200_conv1_fwd(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* w, const float* bias,
                   float* a1, int batch, void* stream);
/* part: scratch [2*batch][320]; dw [9][32]; db [32] */
int b200_conv1_wgrad(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* da1, float* part,
                     float* dw, float* db, int batch, void* stream);
int b200_reduce_partials(float* dst, const float* src, int n, int nparts, int stride, void* stream);
/* Same result as b200_conv1_wgrad, for callers that cannot allocate device memory (their GPU memory
 * belongs to another allocator, or a cudaMalloc's device-wide synchronisation is unaffordable).  This
 * entry point has no counterpart in the reference.  The partial-sum scratch is a 2.5 MiB array built
 * into the library image (held by every GPU that loads the library until the process exits), so no
 * cudaMalloc/cudaFree or allocating routine is called.  batch <= b200_conv1_wgrad_noalloc_max_batch()
 * (1024); larger batches fail with B200_ERR_ARG.  The first call on each device looks the array's
 * address up (cudaGetSymbolAddress, that device current): make that call before any stream capture.
 * All calls on a device share the array and it keeps the last call's contents (nothing needs
 * clearing: every row that is read is rewritten first); calls that can overlap on the GPU race. */
int b200_conv1_wgrad_noalloc_max_batch(void);
int b200_conv1_wgrad_noalloc(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* da1, float* dw,
                             float* db, int batch, void* stream);

/* ---- conv2 on CUDA cores (mnist_ddp.py:52-55; SURVEY N2-N5,N12,N13) ---------------------------- */
int b200_conv2_fwd_simt(const float* a1, const float* w, const float* bias, const void* ctrl, float* pooled,
                        uint8_t* amax, int batch, void* stream);
int b200_conv2_dgrad_simt(const float* dp, const uint8_t* amax, const float* w, const float* a1, float* da1,
                          int batch, void* stream);
int b200_conv2_wgrad_parts(void); /* number of partial slots the wgrad kernels need (18496 floats each) */
int b200_conv2_wgrad_simt(const float* dp, const uint8_t* amax, const float* a1, float* part, float* dw, float* db,
                          int batch, void* stream);
/* ---- fp32 SIMT GEMM (mnist_ddp.py:57; SURVEY N6,N11) ------------------------------------------- */
int b200_sgemm(const float* A, long sam, long sak, const float* B, long sbk, long sbn, float* C, int M, int N, int K,
               int splits, int epi, const float* aux, float scale, void* stream);

/* ---- tcgen05 / TMA / TMEM kernels (SURVEY section 8 rows 3,4,6-9) ------------------------------ */
/* D[M][N] (+split-K partials) = A . B in TF32 with fp32 accumulation.
 *   a_mn = 0: A is [M][K] row-major (K-major operand);  a_mn = 1: A is [K][M] row-major (M-major)
 *   b_mn = 0: B is [N][K] row-major (K-major operand);  b_mn = 1: B is [K][N] row-major (N-major)
 * splits > 1 writes `splits` partial [M][N] matrices back to back.  epi as in b200_sgemm.
 * Host-side set-up (tensor maps) is cached per (pointer, shape). */
int b200_tc_gemm(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn, int splits,
                 int epi, const float* aux, float scale, void* stream);
/* same; fp32_class != 0 selects the 3xTF32 split-operand product (hi.hi + hi.lo + lo.hi, operands split on chip):
 * fp32-class accuracy for the layers the reference runs as fp32 SGEMM (mnist_ddp.py:46,57) */
int b200_tc_gemm_ex(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn, int splits,
                    int epi, const float* aux, float scale, int fp32_class, void* stream);
/* implicit-GEMM conv2 forward from an activation tensor a1[B][26][26][32]: kw taps on the N side (128x192x8 MMAs),
 * fused bias + ReLU + 2x2 max-pool (2-bit argmax) + dropout epilogue; same outputs as _simt */
int b200_conv2_fwd_tc(const float* a1, const float* w, const float* bias, const void* ctrl, float* pooled,
                      uint8_t* amax, int batch, void* stream);
/* conv1 (gather + normalise + conv + bias + ReLU) computed inside that kernel's producer warps: one launch from
 * the uint8 dataset to the pooled activation; a1 is still written (the backward pass reads it).  The engine's
 * forward (mnist_ddp.py:50-55). */
int b200_conv12_fwd_tc(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* w1, const float* b1,
                       float* a1, const float* w2, const float* b2, float* pooled, uint8_t* amax, int batch,
                       void* stream);
/* conv2 backward straight from the pooled gradient dp[B][12][12][64] + argmax codes: the dense dY[B][24][24][64] is
 * never materialised (producer warps expand it into the swizzled operand tiles).  dgrad applies conv1's ReLU mask.
 * wgrad: part = scratch of b200_conv2_wgrad_tc_parts(batch) rows of 18496 floats (dW[9][64][32] then db[64]);
 * dw / db may be NULL when the caller reduces the partial rows itself. */
int b200_conv2_dgrad_tc(const float* dp, const uint8_t* amax, const float* w, const float* a1, float* da1, int batch,
                        void* stream);
int b200_conv2_wgrad_tc_parts(int batch);
int b200_conv2_wgrad_tc(const float* dp, const uint8_t* amax, const float* a1, float* part, float* dw, float* db,
                        int batch, void* stream);

/* ---- classifier head (mnist_ddp.py:57-61,71,97-99; SURVEY N7-N10,N22) -------------------------- */
int b200_head_fused(const float* part, int splits, int batch, const float* b1, const float* w2, const float* b2,
                    const uint8_t* labels, const int32_t* idx, const void* ctrl, float* h, float* dlogits, float* dh,
                    float* loss_i, int* correct_i, float* logp, void* stream);
int b200_head_wgrad(const float* h, const float* dlogits, const float* dh, const float* loss_i, int batch, float* dw2,
                    float* db2, float* db1, float* loss_out, const void* ctrl, float* loss_ring, int ring,
                    float* loss_ring_host /* device alias of mapped pinned host memory, or NULL */, void* stream);
int b200_eval_accumulate(const float* loss_i, const int* correct_i, int batch, double* acc, void* stream);

/* ---- optimizer (mnist_ddp.py:73,176) ------------------------------------------------ */
int b200_adadelta(float* p, const float* g, float* v, float* u, long n, void* ctrl, float rho, float eps,
                  int advance_batch, unsigned int* done_counter, void* stream);

/* ---- training engine (mnist_ddp.py:65-105 train()/test()) ---------------------------- */
typedef struct B200EngineConfig {
  int device;
  int max_batch;        /* largest train or eval batch that will be used */
  int world, rank;
  unsigned long long seed;
  int use_tc;           /* bit mask: 1 conv1+conv2 fwd, 2 fc1 fwd, 4 fc1 bwd, 8 conv2 bwd on tcgen05; 16: fc1 GEMMs as
                           3xTF32 (fp32-class); 0: fp32 CUDA-core kernels everywhere */
  int train_n, test_n;  /* dataset sizes (samples) */
  long param_floats;    /* ParamLayout.total */
  long offsets[8];      /* conv1.w, conv1.b, conv2.w, conv2.b, fc1.w, fc1.b, fc2.w, fc2.b (floats) */
  long bucket_begin[2], bucket_end[2];
  /* optional externally owned (e.g. symmetric) buffers; NULL = the engine allocates them */
  float* params; float* grads[2]; float* sq_avg; float* acc_delta;
} B200EngineConfig;

enum { B200_BUF_PARAMS = 0, B200_BUF_GRADS0, B200_BUF_GRADS1, B200_BUF_SQ_AVG, B200_BUF_ACC_DELTA, B200_BUF_A1,
       B200_BUF_POOLED, B200_BUF_AMAX, B200_BUF_H, B200_BUF_DLOGITS, B200_BUF_DH, B200_BUF_DP, B200_BUF_DA1,
       B200_BUF_LOSS_I, B200_BUF_LOSS_RING, B200_BUF_CTRL, B200_BUF_LOGP, B200_BUF_STAGE_IMG, B200_BUF_STAGE_LBL,
       B200_BUF_FC1_PART, B200_BUF_COUNT };
enum { B200_LOSS_RING = 1024 };

int b200_engine_create(const B200EngineConfig* cfg, void** engine_out);
int b200_engine_destroy(void* engine);
int b200_engine_buffer(void* engine, int which, void** ptr, size_t* bytes);
/* copies a split (0 = train, 1 = test) into HBM; synchronous */
int b200_engine_load_data(void* engine, int split, const uint8_t* images_host, const uint8_t* labels_host, int n);
/* uploads this epoch's sample order (NULL = sequential) and rewinds the cursor; idx_host must stay
 * valid until the stream has consumed it (pinned memory recommended) */
int b200_engine_begin_epoch(void* engine, const int32_t* idx_host, int n, void* stream);
int b200_engine_set_lr(void* engine, float lr, void* stream);
/* on = 0 trains without dropout (for numerics / data-parallel equivalence tests); default on */
int b200_engine_set_dropout(void* engine, int on, void* stream);
/* unfused baseline: with defer = 1 train_step stops after backward; reduce the gradient buffer yourself,
 * then b200_engine_apply_update runs Adadelta and advances the cursor by `advance` samples */
int b200_engine_set_defer_update(void* engine, int defer);
int b200_engine_apply_update(void* engine, int advance, void* stream);
/* one optimizer step on `batch` samples starting at the cursor: forward, backward, (all-reduce,)
 * Adadelta; replayed from a CUDA graph when use_graph != 0.  Asynchronous. */
int b200_engine_train_step(void* engine, int batch, int use_graph, void* stream);
/* same, but the batch comes from host memory: H2D copy of uint8 images/labels, then the step */
int b200_engine_train_step_host(void* engine, const uint8_t* images_host, const uint8_t* labels_host, int batch,
                                int use_graph, void* stream);
/* the data-loader path: gather `batch` samples of a host-resident uint8 dataset by `idx_host` into the engine's
 * pinned staging memory (a 784-byte memcpy per sample instead of the reference's per-sample transform pipeline,
 * mnist_ddp.py:144-168), then as train_step_host.  Every index is checked against `n_samples` (the dataset's size)
 * before anything is read: an index outside [0, n_samples) returns B200_ERR_ARG */
int b200_engine_train_step_gather(void* engine, const uint8_t* images_host, const uint8_t* labels_host,
                                  int n_samples, const int32_t* idx_host, int batch, int use_graph, void* stream);
/* forward only on `batch` samples of a split starting at `first`; accumulates sum-NLL / #correct */
int b200_engine_eval_batch(void* engine, int split, int first, int batch, void* stream);
int b200_engine_eval_reset(void* engine, void* stream);
int b200_engine_eval_result(void* engine, double* nll_sum_host, long long* correct_host, void* stream); /* syncs */
/* mean loss of optimizer step `step` (must be among the last B200_LOSS_RING steps); syncs the stream */
int b200_engine_read_loss(void* engine, unsigned long long step, float* loss_host, void* stream);
int b200_engine_step_count(void* engine, unsigned long long* steps_host);
/* host address of the mapped, pinned loss ring (B200_LOSS_RING floats): the head kernel writes every
 * step's mean loss straight into it (zero-copy D2H); valid once the step's work has completed */
int b200_engine_host_loss_ring(void* engine, float** ring_host);
/* kernels launched per train step at this batch size (for bench.py's gpu_launches) */
int b200_engine_launches_per_step(void* engine, int batch, int* n_host);

/* ---- symmetric memory + fused data-parallel reduction (mnist_ddp.py:173 DistributedDataParallel;
 *      SURVEY N15, N17-N19, section 8 rows 11, 12, 15, 17) ---------------------------------------- */
/* One device allocation per rank ([payload][barrier flags]), mapped into every peer with CUDA IPC. */
int b200_symm_handle_bytes(void);
int b200_symm_flag_bytes(void);
int b200_symm_create(int world, int rank, size_t payload_bytes, void** arena_out);
int b200_symm_export(void* arena, void* handle_out_host);            /* b200_symm_handle_bytes() bytes */
int b200_symm_import(void* arena, const void* all_handles_host);      /* world handles, indexed by rank */
/* adopt memory made symmetric elsewhere (e.g. torch symmetric memory); mc_ptr = NVLS multicast alias or 0 */
int b200_symm_wrap(int world, int rank, size_t payload_bytes, const uint64_t* peer_ptrs_host, uint64_t mc_ptr,
                   void** arena_out);
int b200_symm_has_multicast(void* arena);
int b200_symm_create_emulated(int world, size_t payload_bytes, void** arenas_out); /* W ranks on one GPU */
int b200_symm_peer(void* arena, int rank, void** base_ptr);
int b200_symm_pull(void* arena, int src_rank, size_t float_off, size_t n_floats, void* stream);
int b200_symm_destroy(void* arena);
/* A plan binds a bucket (float offsets inside the payload) to a barrier channel (0..3). */
int b200_comm_plan_create(void* arena, size_t grad_off, size_t param_off, size_t v_off, size_t u_off,
                          size_t n_floats, int channel, const void* ctrl, void** plan_out);
int b200_comm_plan_create_emulated(void** arenas, int world, size_t grad_off, size_t param_off, size_t v_off,
                                   size_t u_off, size_t n_floats, int channel, const void** ctrls, void** plan_out);
int b200_comm_plan_destroy(void* plan);
/* single-GPU test harness of the conv-bucket kernels (conv2_reduce + conv_finalize, csrc/allreduce.cu): arenas with
 * payload [params n][grads n][sq_avg n][acc_delta n][reduced n], n = 19456 (the engine's conv bucket layout) */
int b200_comm_conv_plan_create_emulated(void** arenas, int world, size_t n_floats, const void** ctrls, void** plan_out);
int b200_comm_conv_bucket_emulated(void* plan, int mode, const float* const* c2_parts, int np2,
                                   const float* const* c1_parts, int ng, int advance, unsigned int* counters,
                                   void* stream);
/* text of a pending cross-GPU barrier time-out ("rank r waited for rank q ..."), or "" -- see B200_COMM_TIMEOUT_MS */
const char* b200_comm_fault_message(void);
/* mode 0 one-shot / 1 two-shot / 2 NVLS two-shot (multimem.ld_reduce + multimem.st; needs a multicast arena); opt 1 = fused 1/W + Adadelta, 0 = averaged all-reduce into param_off;
 * token 0 = derive from ctrl->step, else a caller token growing by 2 per call; nblocks 0 = automatic */
int b200_comm_run(void* plan, int mode, int opt, unsigned int token, int nblocks, void* stream);
/* engine created on the arena's buffers ([params][grads][sq_avg][acc_delta][reduced conv gradients],
 * ParamLayout.total floats each); modes: 0 one-shot, 1 two-shot, 2 NVLS, for the fc bucket / the conv2 half */
int b200_ddp_attach(void* engine, void* arena, int mode_bucket0, int mode_bucket1);

#ifdef __cplusplus
}
#endif
#endif /* B200MNIST_H_ */
