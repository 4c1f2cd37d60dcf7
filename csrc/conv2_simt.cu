// conv2 (32->64, 3x3) on CUDA cores: forward, dgrad and wgrad.
//
// These SIMT kernels are (a) the numerics oracle for the tcgen05 implicit-GEMM kernels in
// tc_conv2_fwd.cu / tc_conv2_bwd.cu, and (b) the path used for shapes the tensor-core kernels do not take.  They fuse
// the same epilogues/producers as the tensor-core versions so the surrounding buffers are
// identical:
//   forward : bias + ReLU + 2x2 max-pool (2-bit argmax) + dropout(.25)   (mnist_ddp.py:52-55)
//   dgrad   : un-pool scatter in the loader, ReLU mask of conv1 in the epilogue
//   wgrad   : un-pool in the loader; iterates pooled windows (one non-zero per 2x2 window), so
//             it does a quarter of the dense FLOPs; bias gradient in the same pass
// Layouts: activations NHWC; weights [tap][o][i] (see models/net.py).
#include "common.cuh"

using namespace net;

constexpr int TILE = 8;            // 8x8 output pixels per tile
constexpr int PATCH = TILE + 2;    // 10x10 input pixels
constexpr int W2_ELEMS = 9 * C2 * C1;  // 18432

// ---------------------------------------------------------------------------------------------
// forward.  Persistent blocks; weights live in smem transposed to [tap][i][o] so that a warp's
// 32 lanes (32 consecutive o) read consecutive words.  Thread t: o = t%64, row pair g = t/64
// (rows 2g, 2g+1 of the tile = 4 pooling windows).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 2)
conv2_fwd_simt_kernel(const float* __restrict__ a1, const float* __restrict__ w, const float* __restrict__ bias,
                      const B200Ctrl* __restrict__ ctrl, float* __restrict__ pooled, uint8_t* __restrict__ amax,
                      int batch) {
  extern __shared__ __align__(16) float smem[];
  float* ws = smem;                   // [9][32][64]
  float* patch = smem + W2_ELEMS;     // [100][32]
  for (int e = threadIdx.x; e < W2_ELEMS; e += 256) {
    const int tap = e / (C2 * C1), r = e - tap * (C2 * C1), o = r / C1, i = r - o * C1;
    ws[(tap * C1 + i) * C2 + o] = w[e];
  }
  const int o = threadIdx.x & 63, g = threadIdx.x >> 6;
  const float bo = bias[o];
  const int train = ctrl->train;
  const unsigned long long seed = ctrl->seed, step = ctrl->step;
  const int ntiles = batch * 9;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b = tile / 9, tb = tile - b * 9, ty = (tb / 3) * TILE, tx = (tb % 3) * TILE;
    __syncthreads();  // previous tile's readers done (also covers the weight fill on iteration 0)
    const float* src = a1 + (size_t)b * (H1 * W1 * C1);
    for (int e = threadIdx.x; e < PATCH * PATCH * (C1 / 4); e += 256) {
      const int pix = e >> 3, q = e & 7, py = pix / PATCH, px = pix - py * PATCH;
      reinterpret_cast<float4*>(patch)[e] =
          *reinterpret_cast<const float4*>(src + ((ty + py) * W1 + tx + px) * C1 + q * 4);
    }
    __syncthreads();
    float acc[16];
#pragma unroll
    for (int p = 0; p < 16; ++p) acc[p] = bo;
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      const int kh = tap / 3, kw = tap - kh * 3;
      const float* wt = ws + tap * C1 * C2 + o;
      const float* pt = patch + ((2 * g + kh) * PATCH + kw) * C1;
#pragma unroll 2
      for (int i4 = 0; i4 < C1; i4 += 4) {
        const float w0 = wt[(i4 + 0) * C2], w1 = wt[(i4 + 1) * C2], w2 = wt[(i4 + 2) * C2], w3 = wt[(i4 + 3) * C2];
#pragma unroll
        for (int p = 0; p < 16; ++p) {
          const int r = p >> 3, c = p & 7;
          const float4 v = *reinterpret_cast<const float4*>(pt + (r * PATCH + c) * C1 + i4);
          acc[p] = fmaf(v.x, w0, acc[p]);
          acc[p] = fmaf(v.y, w1, acc[p]);
          acc[p] = fmaf(v.z, w2, acc[p]);
          acc[p] = fmaf(v.w, w3, acc[p]);
        }
      }
    }
    // ReLU + 2x2 max-pool (first maximum wins, like ATen's max_pool2d) + dropout
#pragma unroll
    for (int wdw = 0; wdw < 4; ++wdw) {
      const float v0 = acc[2 * wdw], v1 = acc[2 * wdw + 1], v2 = acc[8 + 2 * wdw], v3 = acc[8 + 2 * wdw + 1];
      float m = v0; int am = 0;
      if (v1 > m) { m = v1; am = 1; }
      if (v2 > m) { m = v2; am = 2; }
      if (v3 > m) { m = v3; am = 3; }
      m = fmaxf(m, 0.f);
      const int ph = (ty >> 1) + g, pw = (tx >> 1) + wdw;
      const uint32_t e = ((uint32_t)(b * HP + ph) * WP + pw) * C2 + o;
      if (train & 2) {
        const uint4 r = dropout_block(seed, step, 1u, e >> 6);
        m = dropout1_keep(r, e & 63u) ? m * (1.0f / (1.0f - P_DROP1)) : 0.f;
      }
      pooled[e] = m;
      amax[e] = (uint8_t)am;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// dY (the gradient at conv2's pre-pool output) is never materialised: element (y, x, o) equals
// dp[y/2][x/2][o] when the stored argmax of that window is (y%2)*2 + x%2, else 0.  `dp` already
// carries the dropout scale and the ReLU mask (applied by the fc1 dgrad epilogue).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float unpool_at(const float* __restrict__ dp, const uint8_t* __restrict__ amax, int b,
                                           int y, int x, int o) {
  if ((unsigned)y >= (unsigned)H2 || (unsigned)x >= (unsigned)W2) return 0.f;
  const size_t e = ((size_t)(b * HP + (y >> 1)) * WP + (x >> 1)) * C2 + o;
  return amax[e] == (uint8_t)(((y & 1) << 1) | (x & 1)) ? dp[e] : 0.f;
}

// dgrad: da1[y][x][i] = relu'(a1) * sum_{kh,kw,o} dY[y-kh][x-kw][o] * W[kh][kw][o][i].
// Tiles of 8x8 over the 26x26 input grid (4x4 tiles, edge tiles partial).  Thread t: i = t%32,
// tile row t/32.  Weights in smem in their native [tap][o][i] order (i across lanes).
__global__ void __launch_bounds__(256, 2)
conv2_dgrad_simt_kernel(const float* __restrict__ dp, const uint8_t* __restrict__ amax, const float* __restrict__ w,
                        const float* __restrict__ a1, float* __restrict__ da1, int batch) {
  extern __shared__ __align__(16) float smem[];
  float* ws = smem;                 // [9][64][32]
  float* patch = smem + W2_ELEMS;   // [100][64]
  for (int e = threadIdx.x; e < W2_ELEMS / 4; e += 256)
    reinterpret_cast<float4*>(ws)[e] = reinterpret_cast<const float4*>(w)[e];
  const int i = threadIdx.x & 31, row = threadIdx.x >> 5;
  const int ntiles = batch * 16;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b = tile >> 4, tb = tile & 15, ty = (tb >> 2) * TILE, tx = (tb & 3) * TILE;
    __syncthreads();
    for (int e = threadIdx.x; e < PATCH * PATCH * C2; e += 256) {
      const int pix = e >> 6, o = e & 63, py = pix / PATCH, px = pix - py * PATCH;
      patch[e] = unpool_at(dp, amax, b, ty - 2 + py, tx - 2 + px, o);
    }
    __syncthreads();
    float acc[TILE];
#pragma unroll
    for (int p = 0; p < TILE; ++p) acc[p] = 0.f;
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      const int kh = tap / 3, kw = tap - kh * 3;
      const float* wt = ws + tap * C2 * C1 + i;
      const float* pt = patch + ((row + 2 - kh) * PATCH + (2 - kw)) * C2;
#pragma unroll 2
      for (int o4 = 0; o4 < C2; o4 += 4) {
        const float w0 = wt[(o4 + 0) * C1], w1 = wt[(o4 + 1) * C1], w2 = wt[(o4 + 2) * C1], w3 = wt[(o4 + 3) * C1];
#pragma unroll
        for (int p = 0; p < TILE; ++p) {
          const float4 v = *reinterpret_cast<const float4*>(pt + p * C2 + o4);
          acc[p] = fmaf(v.x, w0, acc[p]);
          acc[p] = fmaf(v.y, w1, acc[p]);
          acc[p] = fmaf(v.z, w2, acc[p]);
          acc[p] = fmaf(v.w, w3, acc[p]);
        }
      }
    }
    const int y = ty + row;
    if (y < H1) {
#pragma unroll
      for (int p = 0; p < TILE; ++p) {
        const int x = tx + p;
        if (x < W1) {
          const size_t e = ((size_t)(b * H1 + y) * W1 + x) * C1 + i;
          da1[e] = a1[e] > 0.f ? acc[p] : 0.f;
        }
      }
    }
  }
}

// wgrad: dW[tap][o][i] = sum_{b,y,x} dY[b][y][x][o] * a1[b][y+kh][x+kw][i]; db[o] = sum dY.
// Thread t: o = t%64, i-octet ig = t/64; 72 accumulators.  Per pooled window only the argmax
// pixel contributes.  Each block writes one partial [18432 + 64]; a fixed-order reduce follows.
constexpr int CONV2_PART = W2_ELEMS + C2;
__global__ void __launch_bounds__(256, 1)
conv2_wgrad_simt_kernel(const float* __restrict__ dp, const uint8_t* __restrict__ amax,
                        const float* __restrict__ a1, float* __restrict__ part, int batch) {
  __shared__ __align__(16) float patch[PATCH * PATCH * C1];  // 12.8 KB
  __shared__ float dwin[16][C2];                              // pooled grads of the 4x4 windows
  __shared__ uint8_t awin[16][C2];
  const int o = threadIdx.x & 63, ig = threadIdx.x >> 6;
  float acc[9][8];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[t][j] = 0.f;
  float accb = 0.f;
  const int ntiles = batch * 9;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b = tile / 9, tb = tile - b * 9, ty = (tb / 3) * TILE, tx = (tb % 3) * TILE;
    __syncthreads();
    const float* src = a1 + (size_t)b * (H1 * W1 * C1);
    for (int e = threadIdx.x; e < PATCH * PATCH * (C1 / 4); e += 256) {
      const int pix = e >> 3, q = e & 7, py = pix / PATCH, px = pix - py * PATCH;
      reinterpret_cast<float4*>(patch)[e] =
          *reinterpret_cast<const float4*>(src + ((ty + py) * W1 + tx + px) * C1 + q * 4);
    }
    for (int e = threadIdx.x; e < 16 * C2; e += 256) {
      const int wdw = e >> 6, oo = e & 63;
      const size_t g = ((size_t)(b * HP + (ty >> 1) + (wdw >> 2)) * WP + (tx >> 1) + (wdw & 3)) * C2 + oo;
      dwin[wdw][oo] = dp[g];
      awin[wdw][oo] = amax[g];
    }
    __syncthreads();
#pragma unroll 1
    for (int wdw = 0; wdw < 16; ++wdw) {
      const float d = dwin[wdw][o];
      const int a = awin[wdw][o];
      const int ly = ((wdw >> 2) << 1) + (a >> 1), lx = ((wdw & 3) << 1) + (a & 1);
      accb += d;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int kh = tap / 3, kw = tap - kh * 3;
        const float* pt = patch + ((ly + kh) * PATCH + lx + kw) * C1 + ig * 8;
        const float4 v0 = *reinterpret_cast<const float4*>(pt), v1 = *reinterpret_cast<const float4*>(pt + 4);
        acc[tap][0] = fmaf(d, v0.x, acc[tap][0]); acc[tap][1] = fmaf(d, v0.y, acc[tap][1]);
        acc[tap][2] = fmaf(d, v0.z, acc[tap][2]); acc[tap][3] = fmaf(d, v0.w, acc[tap][3]);
        acc[tap][4] = fmaf(d, v1.x, acc[tap][4]); acc[tap][5] = fmaf(d, v1.y, acc[tap][5]);
        acc[tap][6] = fmaf(d, v1.z, acc[tap][6]); acc[tap][7] = fmaf(d, v1.w, acc[tap][7]);
      }
    }
  }
  float* dst = part + (size_t)blockIdx.x * CONV2_PART;
#pragma unroll
  for (int tap = 0; tap < 9; ++tap) {
    float* q = dst + (tap * C2 + o) * C1 + ig * 8;
    *reinterpret_cast<float4*>(q) = make_float4(acc[tap][0], acc[tap][1], acc[tap][2], acc[tap][3]);
    *reinterpret_cast<float4*>(q + 4) = make_float4(acc[tap][4], acc[tap][5], acc[tap][6], acc[tap][7]);
  }
  if (ig == 0) dst[W2_ELEMS + o] = accb;
}

static int sm_count() { return b200_sm_count(); }

extern "C" int b200_conv2_fwd_simt(const float* a1, const float* w, const float* bias, const void* ctrl,
                                   float* pooled, uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_fwd_simt: batch must be > 0");
  const int smem = (W2_ELEMS + PATCH * PATCH * C1) * 4;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr = true;
  }
  const int grid = min(batch * 9, 2 * sm_count());
  conv2_fwd_simt_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(a1, w, bias, (const B200Ctrl*)ctrl, pooled, amax, batch);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv2_dgrad_simt(const float* dp, const uint8_t* amax, const float* w, const float* a1,
                                     float* da1, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_dgrad_simt: batch must be > 0");
  const int smem = (W2_ELEMS + PATCH * PATCH * C2) * 4;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr = true;
  }
  const int grid = min(batch * 16, 2 * sm_count());
  conv2_dgrad_simt_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(dp, amax, w, a1, da1, batch);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv2_wgrad_parts(void) { return sm_count(); }

extern "C" int b200_conv2_wgrad_simt(const float* dp, const uint8_t* amax, const float* a1, float* part,
                                     float* dw, float* db, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_wgrad_simt: batch must be > 0");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = min(batch * 9, sm_count());
  conv2_wgrad_simt_kernel<<<grid, 256, 0, s>>>(dp, amax, a1, part, batch);
  b200_count_launch();
  if (dw) b200_launch_reduce_partials(dw, part, W2_ELEMS, grid, CONV2_PART, s);
  if (db) b200_launch_reduce_partials(db, part + W2_ELEMS, C2, grid, CONV2_PART, s);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
