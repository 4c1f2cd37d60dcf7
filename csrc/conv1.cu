// conv1: 1->32 channels, 3x3, stride 1, no padding (mnist_ddp.py:42,50-51).
//
// K = 9 is far below any tcgen05 tile, so both directions are SIMT kernels whose job is to move
// bytes at HBM speed:
//   forward : batch gather (index lookup into the HBM-resident uint8 dataset) + ToTensor/Normalize
//             ((x/255 - .1307)/.3081, mnist_ddp.py:153-156) + conv + bias + ReLU, one pass, NHWC out.
//             Replaces the reference's DataLoader worker, H2D copy, cuDNN conv, bias add and ReLU
//             kernels.
//             The training step and evaluation run this inside the conv2 forward kernel (tc_conv2_fwd.cu, producer
//             warps); the stand-alone kernel remains for the fp32 CUDA-core path (B200_TC bit 0 cleared) and as oracle.
//   wgrad   : dW[tap][o] = sum da1 * x, db[o] = sum da1; input needs no gradient.
#include <mutex>

#include "common.cuh"

using namespace net;

__device__ __forceinline__ float normalize_px(uint8_t v) {
  // same operation order as ToTensor (div 255) then Normalize (sub mean, div std)
  return __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), MEAN), STD);
}

__device__ __forceinline__ int sample_of(const int32_t* __restrict__ idx, const B200Ctrl* __restrict__ ctrl, int i) {
  int pos = ctrl->pos + i;
  return idx ? idx[pos] : pos;
}

// grid = batch, block = 256.  Thread t: channel quad t%8, pixels t/8 + 32*k.
__global__ void __launch_bounds__(256)
conv1_fwd_kernel(const uint8_t* __restrict__ images, const int32_t* __restrict__ idx,
                 const B200Ctrl* __restrict__ ctrl, const float* __restrict__ w,  // [9][32]
                 const float* __restrict__ bias, float* __restrict__ a1) {       // [B][26][26][32]
  __shared__ float x[IMG * IMG];
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  const int sample = sample_of(idx, ctrl, b);
  const uint8_t* src = images + (size_t)sample * (IMG * IMG);
  if (threadIdx.x < IMG * IMG / 4) {
    uchar4 v = reinterpret_cast<const uchar4*>(src)[threadIdx.x];
    x[threadIdx.x * 4 + 0] = normalize_px(v.x);
    x[threadIdx.x * 4 + 1] = normalize_px(v.y);
    x[threadIdx.x * 4 + 2] = normalize_px(v.z);
    x[threadIdx.x * 4 + 3] = normalize_px(v.w);
  }
  const int cq = threadIdx.x & 7;
  float wr[9][4], br[4];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    float4 v = *reinterpret_cast<const float4*>(w + t * C1 + cq * 4);
    wr[t][0] = v.x; wr[t][1] = v.y; wr[t][2] = v.z; wr[t][3] = v.w;
  }
  {
    float4 v = *reinterpret_cast<const float4*>(bias + cq * 4);
    br[0] = v.x; br[1] = v.y; br[2] = v.z; br[3] = v.w;
  }
  __syncthreads();
  float* out = a1 + (size_t)b * (H1 * W1 * C1);
  for (int p = threadIdx.x >> 3; p < H1 * W1; p += 32) {
    const int y = p / W1, xx = p - y * W1;
    float acc[4] = {br[0], br[1], br[2], br[3]};
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const float v = x[(y + kh) * IMG + xx + kw];
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[c] = fmaf(v, wr[kh * 3 + kw][c], acc[c]);
      }
    float4 o = make_float4(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
    *reinterpret_cast<float4*>(out + p * C1 + cq * 4) = o;
  }
}

// grid = 2*batch, block = 256 (8 pixel groups x 32 channels).  Writes one partial row per block:
// part[blk][0..287] = dW[tap][o], part[blk][288..319] = db[o].  A fixed-order reduction follows.
constexpr int CONV1_PART = 320;
__global__ void __launch_bounds__(256)
conv1_wgrad_kernel(const uint8_t* __restrict__ images, const int32_t* __restrict__ idx,
                   const B200Ctrl* __restrict__ ctrl, const float* __restrict__ da1,  // [B][26][26][32]
                   float* __restrict__ part, float* __restrict__ gpart = nullptr,
                   unsigned int* __restrict__ counters = nullptr) {
  // x (the image) and red (the cross-warp reduction) share storage: 10 KB of static shared memory lets
  // these CTAs co-reside with a conv2 wgrad CTA (214 KB) that runs on the engine's branch stream
  __shared__ __align__(16) float sh[8 * 10 * C1];
  float* x = sh;
  float (*red)[10][C1] = reinterpret_cast<float (*)[10][C1]>(sh);
  pdl_trigger();
  pdl_wait();
  // two blocks per image (13 output rows each) so that 2*batch blocks keep every SM busy
  const int b = blockIdx.x >> 1, half = blockIdx.x & 1;
  const int sample = sample_of(idx, ctrl, b);
  const uint8_t* src = images + (size_t)sample * (IMG * IMG);
  if (threadIdx.x < IMG * IMG / 4) {
    uchar4 v = reinterpret_cast<const uchar4*>(src)[threadIdx.x];
    x[threadIdx.x * 4 + 0] = normalize_px(v.x);
    x[threadIdx.x * 4 + 1] = normalize_px(v.y);
    x[threadIdx.x * 4 + 2] = normalize_px(v.z);
    x[threadIdx.x * 4 + 3] = normalize_px(v.w);
  }
  __syncthreads();
  const int o = threadIdx.x & 31, g = threadIdx.x >> 5;
  const float* in = da1 + (size_t)b * (H1 * W1 * C1);
  float acc[10];
#pragma unroll
  for (int t = 0; t < 10; ++t) acc[t] = 0.f;
  constexpr int HALF = (H1 / 2) * W1;  // 338 pixels
  const int p0 = half * HALF;
  for (int pb = g; pb < HALF; pb += 64) {
    float d[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) d[u] = (pb + 8 * u < HALF) ? in[(p0 + pb + 8 * u) * C1 + o] : 0.f;  // 8 loads in flight
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int p = min(p0 + pb + 8 * u, H1 * W1 - 1);
      const int y = p / W1, xx = p - y * W1;
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) acc[kh * 3 + kw] = fmaf(d[u], x[(y + kh) * IMG + xx + kw], acc[kh * 3 + kw]);
      acc[9] += d[u];
    }
  }
  __syncthreads();  // everybody is done with x
#pragma unroll
  for (int t = 0; t < 10; ++t) red[g][t][o] = acc[t];
  __syncthreads();
  for (int e = threadIdx.x; e < 10 * C1; e += 256) {
    const int t = e / C1, oo = e - t * C1;
    float s = 0.f;
#pragma unroll
    for (int gg = 0; gg < 8; ++gg) s += red[gg][t][oo];
    part[(size_t)blockIdx.x * CONV1_PART + e] = s;  // e = t*32+o: taps first (288), then bias (32)
  }
  if (gpart) group_reduce_rows(part, CONV1_PART, CONV1_PART, gpart, counters);
}

// dst[e] = sum_{p < nparts} src[p*stride + e]  — fixed order, so results are run-to-run identical.
// block = 64 outputs x 16 part-slices: slice g sums parts g, g+16, ... serially, slices are combined
// in a fixed order through shared memory.
__global__ void __launch_bounds__(1024)
reduce_partials_kernel(float* __restrict__ dst, const float* __restrict__ src, int n, int nparts, int stride) {
  __shared__ float red[16][64];
  const int j = threadIdx.x & 63, g = threadIdx.x >> 6;
  const int e = blockIdx.x * 64 + j;
  float s = 0.f;
  if (e < n)
    for (int p = g; p < nparts; p += 16) s += src[(size_t)p * stride + e];
  red[g][j] = s;
  __syncthreads();
  if (g == 0 && e < n) {
    float t = 0.f;
#pragma unroll
    for (int q = 0; q < 16; ++q) t += red[q][j];
    dst[e] = t;
  }
}

void b200_launch_reduce_partials(float* dst, const float* src, int n, int nparts, int stride, cudaStream_t s) {
  reduce_partials_kernel<<<ceil_div(n, 64), 1024, 0, s>>>(dst, src, n, nparts, stride);
  b200_count_launch();
}

extern "C" int b200_conv1_fwd(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* w,
                              const float* bias, float* a1, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv1_fwd: batch must be > 0");
  B200_CHECK(b200_launch_serial(conv1_fwd_kernel, dim3(batch), dim3(256), 0, (cudaStream_t)stream, images, idx,
                         (const B200Ctrl*)ctrl, w, bias, a1));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv1_wgrad(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* da1,
                                float* part, float* dw, float* db, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv1_wgrad: batch must be > 0");
  cudaStream_t s = (cudaStream_t)stream;
  B200_CHECK(b200_launch(conv1_wgrad_kernel, dim3(2 * batch), dim3(256), 0, s, images, idx, (const B200Ctrl*)ctrl, da1, part,
                         (float*)nullptr, (unsigned int*)nullptr));
  b200_count_launch();
  if (dw) b200_launch_reduce_partials(dw, part, 288, 2 * batch, CONV1_PART, s);  // NULL: caller reduces the partials
  if (db) b200_launch_reduce_partials(db, part + 288, 32, 2 * batch, CONV1_PART, s);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int b200_conv1_wgrad_grouped(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* da1, float* part,
                             float* gpart, unsigned int* counters, int batch, cudaStream_t s) {
  B200_REQUIRE(batch > 0 && gpart && counters, "conv1_wgrad_grouped: bad arguments");
  B200_CHECK(b200_launch(conv1_wgrad_kernel, dim3(2 * batch), dim3(256), 0, s, images, idx, (const B200Ctrl*)ctrl, da1, part,
                         gpart, counters));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_reduce_partials(float* dst, const float* src, int n, int nparts, int stride, void* stream) {
  b200_launch_reduce_partials(dst, src, n, nparts, stride, (cudaStream_t)stream);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// conv1 weight gradient for callers that must not allocate device memory.
//
// b200_conv1_wgrad needs a scratch buffer of 2*batch*320 floats for its per-block partial sums,
// and somebody has to allocate it.  Some users of this library cannot: a host process whose GPU
// memory has been handed over in full to another allocator (a framework memory pool sized to the
// whole device leaves nothing for a stray cudaMalloc), or a control loop with a latency budget
// that cudaMalloc/cudaFree would break because both synchronise the whole device.  For them the
// scratch lives in the library image itself: a file-scope device array that every GPU loading
// this library carries until the process exits (B200_NOALLOC_MAX_BATCH * 2 * 320 floats
// = 2.5 MiB), so the entry point below never calls cudaMalloc, cudaFree, or any routine that
// allocates.
//
// Things a caller has to know:
//  * batch may not exceed B200_NOALLOC_MAX_BATCH (1024); larger batches are refused with
//    B200_ERR_ARG and a message, nothing is launched.
//  * Inside host code the name `g_noalloc_scratch` is only a handle for the symbol, not a device
//    address; the real address is looked up with cudaGetSymbolAddress the first time the entry
//    point runs on a given device (with that device current) and remembered per device.  That
//    first call per device does the look-up, so make it before starting any stream capture.
//    The look-up is for the CURRENT device: `stream` must belong to it (a stream of another device
//    fails with an invalid-resource-handle launch error).
//  * All calls on one device share the one array, and it keeps whatever the previous call left
//    in it.  Nothing in it has to start at zero: a call overwrites every partial row it later
//    reads (rows 0 .. 2*batch-1) before the reduction kernels read them.  Two calls whose kernels
//    can overlap on the GPU (different streams without ordering between them) would write the
//    same rows at the same time and corrupt each other: keep calls on one device ordered.
// ---------------------------------------------------------------------------------------------
#define B200_NOALLOC_MAX_BATCH 1024
static __device__ __align__(256) float g_noalloc_scratch[(size_t)B200_NOALLOC_MAX_BATCH * 2 * CONV1_PART];

static int noalloc_scratch_for_current_device(float** out) {
  constexpr int MAX_DEVICES = 64;
  static float* cached[MAX_DEVICES] = {nullptr};
  static std::mutex mu;
  int dev = -1;
  B200_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "conv1_wgrad_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lk(mu);
  if (!cached[dev]) {
    float* p = nullptr;
    B200_CHECK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));  // device `dev` is current here
    B200_REQUIRE(p != nullptr, "conv1_wgrad_noalloc: symbol look-up returned a null address on device %d", dev);
    cached[dev] = p;
  }
  *out = cached[dev];
  return B200_OK;
}

extern "C" int b200_conv1_wgrad_noalloc_max_batch(void) { return B200_NOALLOC_MAX_BATCH; }

extern "C" int b200_conv1_wgrad_noalloc(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* da1,
                                        float* dw, float* db, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv1_wgrad_noalloc: batch must be > 0");
  // idx may be NULL (sequential addressing); everything else is dereferenced on the device
  B200_REQUIRE(images && ctrl && da1 && dw && db,
               "conv1_wgrad_noalloc: images, ctrl, da1, dw and db must all be non-NULL device pointers");
  B200_REQUIRE(batch <= B200_NOALLOC_MAX_BATCH,
               "conv1_wgrad_noalloc: batch %d exceeds the built-in scratch (max %d samples); use b200_conv1_wgrad "
               "with a caller-provided scratch buffer instead", batch, B200_NOALLOC_MAX_BATCH);
  float* scratch = nullptr;
  int r = noalloc_scratch_for_current_device(&scratch);
  if (r != B200_OK) return r;
  cudaStream_t s = (cudaStream_t)stream;
  conv1_wgrad_kernel<<<2 * batch, 256, 0, s>>>(images, idx, (const B200Ctrl*)ctrl, da1, scratch, nullptr, nullptr);
  b200_count_launch();
  b200_launch_reduce_partials(dw, scratch, 288, 2 * batch, CONV1_PART, s);
  b200_launch_reduce_partials(db, scratch + 288, 32, 2 * batch, CONV1_PART, s);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
