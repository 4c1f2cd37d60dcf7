// Shared device/host helpers for libb200mnist (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/b200mnist.h"

#define B200_CHECK(expr)                                                                  \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      b200_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return B200_ERR_CUDA;                                                               \
    }                                                                                     \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CHECK(cudaGetLastError())

#define B200_REQUIRE(cond, ...)        \
  do {                                 \
    if (!(cond)) {                     \
      b200_set_error(__VA_ARGS__);     \
      return B200_ERR_ARG;             \
    }                                  \
  } while (0)

void b200_set_error(const char* fmt, ...);
void b200_count_launch(int n = 1);

// What a cross-GPU barrier writes into mapped pinned HOST memory when it gives up waiting for a peer (allreduce.cu),
// just before it traps.  The trap surfaces on the host as "unspecified launch failure"; b200_set_error() looks at the
// registered records and appends which rank waited for which peer, so the message names the culprit.
struct CommFault {
  unsigned int valid;  // 0 empty, 1 being written, 2 complete
  int rank, world, channel, kind /* 0 entry barrier, 1 exit barrier, 2 chunk owner waiting for contributions */, block, peer;
  unsigned int expected, seen;
  unsigned long long step, waited_ns;
};
void b200_register_fault_record(CommFault* host_ptr);
void b200_unregister_fault_record(CommFault* host_ptr);
// dst[e] = sum_p src[p*stride + e], fixed order (conv1.cu)
void b200_launch_reduce_partials(float* dst, const float* src, int n, int nparts, int stride, cudaStream_t s);
int b200_sm_count();

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------
// Every kernel of the training step calls pdl_trigger() first (the NEXT kernel of the stream / graph may
// start being scheduled as soon as all CTAs of this one have started) and pdl_wait() before its first
// global-memory access (blocks until the PREVIOUS kernel has completed and its writes are visible).
// Launch latency, block scheduling and the on-chip prologue (barrier init, TMEM allocation) of kernel
// N+1 thereby overlap the tail of kernel N.  Kernels launched without the attribute behave as before.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
bool b200_pdl_enabled(const void* kernel);  // runtime.cu: experimental, only with B200_PDL=1 (see the note there)

// Every kernel of the step asks for the maximum shared-memory carve-out, also those that use little or
// no shared memory.  An SM only runs CTAs of kernels with the same L1/shared split at the same time, and
// the persistent tcgen05 kernels need (nearly) all 228 KB: with the default preference a small kernel on
// a branch stream would wait for those CTAs to drain instead of running beside them (measured).
void b200_prefer_max_shared(const void* kernel);  // runtime.cu; once per kernel

template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch_ex(bool allow_pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                  cudaStream_t s, Args&&... args) {
  b200_prefer_max_shared((const void*)kernel);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (allow_pdl && b200_pdl_enabled((const void*)kernel)) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                               Args&&... args) {
  return b200_launch_ex(true, kernel, grid, block, smem, s, static_cast<Args&&>(args)...);
}
// for the first kernel of a step: it follows copies, other streams' joins and the previous step's end
// marker, and is always launched with full stream serialisation
template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch_serial(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                      Args&&... args) {
  return b200_launch_ex(false, kernel, grid, block, smem, s, static_cast<Args&&>(args)...);
}

// ---- second-level reduction of per-CTA partial rows, inside the producing kernel ---------------------
// Kernels that leave one partial row per CTA (conv1 wgrad: 2*batch rows) call this
// at their end: of every GROUP_ROWS consecutive CTAs the one that finishes last adds the group's rows
// (fixed order, so still deterministic) into gpart[group].  The step's final kernel then has ~8x fewer
// rows to walk through, which is what its latency is made of.  counters: one zero-initialised word per
// group; the combining CTA resets it, so the kernel can be replayed from a graph.
#ifdef __CUDACC__
constexpr int GROUP_ROWS = 8;
__device__ __forceinline__ void group_reduce_rows(const float* part, int stride, int n, float* gpart,
                                                  unsigned int* counters) {
  __shared__ int s_last;
  const int grp = blockIdx.x / GROUP_ROWS;
  const int gsize = min(GROUP_ROWS, (int)gridDim.x - grp * GROUP_ROWS);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(&counters[grp], 1u);
    s_last = prev == (unsigned int)(gsize - 1);
    if (s_last) counters[grp] = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const float* src = part + (size_t)grp * GROUP_ROWS * stride;
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    float v[GROUP_ROWS];
#pragma unroll
    for (int q = 0; q < GROUP_ROWS; ++q) v[q] = q < gsize ? __ldcg(src + (size_t)q * stride + e) : 0.f;
    float s = v[0];
#pragma unroll
    for (int q = 1; q < GROUP_ROWS; ++q) s += v[q];
    gpart[(size_t)grp * stride + e] = s;
  }
}
#endif
// engine-internal variant of b200_conv1_wgrad that also combines groups of partial rows
int b200_conv1_wgrad_grouped(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* da1, float* part,
                             float* gpart, unsigned int* counters, int batch, cudaStream_t s);
inline int b200_group_count(int rows) { return (rows + 7) / 8; }

// ---- model constants (mnist_ddp.py:42-47) -------------------------------------------------
namespace net {
constexpr int IMG = 28;
constexpr int C1 = 32, H1 = 26, W1 = 26;   // conv1 output (NHWC)
constexpr int C2 = 64, H2 = 24, W2 = 24;   // conv2 output before pooling
constexpr int HP = 12, WP = 12;            // pooled
constexpr int FC1_IN = HP * WP * C2;       // 9216, (h, w, c) order internally
constexpr int FC1_OUT = 128;
constexpr int NCLS = 10;
constexpr float P_DROP1 = 0.25f, P_DROP2 = 0.5f;
constexpr float MEAN = 0.1307f, STD = 0.3081f;
}  // namespace net

// ---- device control block: everything a replayed CUDA graph needs to read ---------------
// One per engine, in device memory.  Kernels read it; the last kernel of a step advances it.
struct B200Ctrl {
  unsigned long long step;   // global optimizer step, also the dropout Philox offset
  unsigned long long seed;   // dropout seed (same on every rank, mnist_ddp.py:140)
  int pos;                   // first sample of this step inside the epoch's index array
  int train;                 // bit 0: training step (backward seed is produced); bit 1: dropout active
  float lr;                  // Adadelta lr for this epoch (StepLR applied on the host)
  float inv_world;           // 1 / world_size
  unsigned int epoch_flag;   // monotonically increasing cross-GPU barrier epoch
  int reserved[3];
};

// ---- Philox4x32-10 (counter-based RNG for dropout) -------------------------------------------
__host__ __device__ __forceinline__ uint32_t b200_mulhi(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

__host__ __device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = b200_mulhi(M0, c.x), lo0 = M0 * c.x;
    uint32_t hi1 = b200_mulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

// Dropout masks are a pure function of (seed, optimizer step, layer, element index), so any kernel -- CUDA-core or
// tensor-core, forward or a test -- regenerates the same mask without storing it.  One Philox call yields 128 bits
// for a BLOCK of consecutive elements:
//   layer 1 (p = .25): 64 elements per block, 2 bits each, keep iff the two bits are not both zero  (P = 3/4 exactly)
//   layer 2 (p = .5) : 128 elements per block, 1 bit each, keep iff the bit is set
// A pooled pixel has 64 channels and a sample 128 hidden units, so a block is exactly one pixel / one sample: the
// warp that owns it needs ONE Philox evaluation (the 10-round Philox was a third of the conv2 epilogue's
// instructions when every 4 elements had their own call).
__host__ __device__ __forceinline__ uint4 dropout_block(unsigned long long seed, unsigned long long step,
                                                        uint32_t layer, uint32_t block) {
  uint4 ctr = make_uint4(block, layer, (uint32_t)step, (uint32_t)(step >> 32));
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  return philox4x32_10(ctr, key);
}
__host__ __device__ __forceinline__ uint32_t block_word(const uint4& r, uint32_t w) {
  return w == 0 ? r.x : w == 1 ? r.y : w == 2 ? r.z : r.w;
}
// layer 1: keep flag of element `i` (0..63) of its block
__host__ __device__ __forceinline__ bool dropout1_keep(const uint4& r, uint32_t i) {
  return ((block_word(r, i >> 4) >> (2 * (i & 15))) & 3u) != 0u;
}
// layer 2: keep flag of element `i` (0..127) of its block
__host__ __device__ __forceinline__ bool dropout2_keep(const uint4& r, uint32_t i) {
  return ((block_word(r, i >> 5) >> (i & 31)) & 1u) != 0u;
}

// ---- small device utilities -----------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float4 ld_nc_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

// One Adadelta element update (LIB:torch/optim/adadelta.py:292-302), shared by the single-GPU
// kernel (adadelta.cu) and the fused all-reduce kernels (allreduce.cu).
__device__ __forceinline__ void adadelta_update(float g, float& p, float& v, float& u, float lr, float rho, float eps) {
  v = fmaf(v, rho, (1.f - rho) * g * g);
  const float d = __fdiv_rn(sqrtf(u + eps), sqrtf(v + eps)) * g;
  u = fmaf(u, rho, (1.f - rho) * d * d);
  p = fmaf(-lr, d, p);
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
