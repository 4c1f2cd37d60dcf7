// conv2 dgrad, fourth generation (DRAFT: compiles, NOT yet run on a GPU, not used by default).
//
// Same MMA formulation as tc_conv2_v3.cu (kw-scatter, N = 96), but the dense dY tensor is never read:
// four producer warps expand the pooled gradient dp[B][12][12][64] and the 2-bit arg-max codes straight
// into the shared-memory patch that the v3 kernel gets from TMA (max-pool backward = "the value goes to
// the arg-max position of its 2x2 window, zeros elsewhere").  That takes unpool_dense_kernel (7-13 us at
// batch 200, 604 MB of writes at batch 4096) off the critical path: with this kernel the un-pool is only
// needed by conv2 wgrad and moves to that kernel's branch stream (engine.cu, B200_TC bit 64).
//
// The patch must look exactly as TMA with CU_TENSOR_MAP_SWIZZLE_128B would have written it, because the
// tensor core applies the same XOR on absolute shared-memory address bits when it reads the operand
// (DESIGN.md section 3.3): a pixel is one 128-byte row (32 channels of the chunk), and inside each
// 1024-byte group of 8 pixels the 16-byte chunk index is XORed with (pixel & 7):
//     byte offset of (row r, pixel x, channel quad cq) = r*4096 + x*128 + ((cq ^ (x & 7)) << 4)
// Columns 24..31 of every patch row are zero (written once at kernel start); rows outside the image are
// written as zeros for every item.  Generic-proxy stores are made visible to the tensor core with
// fence.proxy.async before the producer warp arrives on the "patch full" barrier.
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "tc_common.cuh"

using namespace tc;
using namespace net;

namespace {

constexpr int TAP_BYTES = 64 * 128;
constexpr int P_ROWS = 18, P_PITCH = 32;
constexpr int ROW_BYTES = P_PITCH * 128;
constexpr int PATCH_BYTES = P_ROWS * ROW_BYTES;
constexpr int DG4_SMEM = 9 * TAP_BYTES + 2 * PATCH_BYTES + 1024;
constexpr int EPI_WARPS = 8, PROD_WARPS = 4;
constexpr int DG4_THREADS = (EPI_WARPS + PROD_WARPS + 1) * 32;  // 416

__global__ void __launch_bounds__(DG4_THREADS, 1)
conv2_dgrad_tc4_kernel(const __grid_constant__ CUtensorMap mapW, const float* __restrict__ dp,
                       const uint8_t* __restrict__ amax, const float* __restrict__ a1, float* __restrict__ da1,
                       int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                     // [kh][kw][64 o][32 i], 32B-atom swizzle: N-major B operand
  uint8_t* patch = smem + 9 * TAP_BYTES;   // [2 channel chunks][18][32][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[4], tempty[4];
  __shared__ uint32_t tmem_base_smem;

  // warps 0..7 epilogue (id % 4 = TMEM sub-partition), 8..11 producers, 12 MMA issuer
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_epi = warp_id < EPI_WARPS, is_mma = warp_id == EPI_WARPS + PROD_WARPS;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&pfull[i], PROD_WARPS); mbar_init(&pempty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], EPI_WARPS); }
    fence_barrier_init();
  }
  // the pad columns (pixels 24..31) are never written again: clear the whole patch area once
  for (int o = threadIdx.x * 16; o < 2 * PATCH_BYTES; o += DG4_THREADS * 16)
    *reinterpret_cast<float4*>(patch + o) = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_proxy_async();
  if (is_mma) tmem_alloc(&tmem_base_smem, 512);  // 4 accumulator slots of 128 columns (96 used)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  // neighbours swapped in odd rounds so that every CTA sees both 4-tile and 3-tile items (see v3)
  const bool flip = (gridDim.x & 1) == 0;
  auto item_of = [&](int it) { return (int)(blockIdx.x + it * gridDim.x) ^ ((flip && (it & 1)) ? 1 : 0); };

  if (!is_epi && !is_mma) {
    // ---- producers: weights by TMA (once), dY patches by expansion of (dp, amax)
    const int ptid = threadIdx.x - EPI_WARPS * 32;  // 0..127
    if (ptid < 32) {
      const bool leader = elect_one();
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t)
        if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
    }
    for (int it = 0; it < my_items; ++it) {
      const int item = item_of(it), n = item >> 1, y0 = (item & 1) * 16;
      const int ph0 = (y0 - 2) / 2;  // -1 or 7: pooled row of patch rows 0 and 1
      for (int kc = 0; kc < 2; ++kc) {
        mbar_wait(&pempty[kc], (it & 1) ^ 1);
        uint8_t* pb = patch + kc * PATCH_BYTES;
        // work unit = (pooled row k of the patch, pooled column pw, channel quad cq): 9 x 12 x 8
        for (int u = ptid; u < 9 * WP * 8; u += PROD_WARPS * 32) {
          const int cq = u & 7, pw = (u >> 3) % WP, k = u / (8 * WP);
          const int ph = ph0 + k;
          float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
          uchar4 a = make_uchar4(0, 0, 0, 0);
          if (ph >= 0 && ph < HP) {
            const size_t e = (((size_t)n * HP + ph) * WP + pw) * C2 + kc * 32 + cq * 4;
            d = *reinterpret_cast<const float4*>(dp + e);
            a = *reinterpret_cast<const uchar4*>(amax + e);
          }
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const int r = 2 * k + (s >> 1), x = 2 * pw + (s & 1);
            const float4 v = make_float4(a.x == s ? d.x : 0.f, a.y == s ? d.y : 0.f, a.z == s ? d.z : 0.f,
                                         a.w == s ? d.w : 0.f);
            *reinterpret_cast<float4*>(pb + r * ROW_BYTES + x * 128 + ((cq ^ (x & 7)) << 4)) = v;
          }
        }
        fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's operand reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&pfull[kc]);
      }
    }
  } else if (is_mma) {
    const bool leader = elect_one();
    constexpr uint32_t idesc = idesc_tf32(128, 96, 0, 1);
    mbar_wait(&w_bar, 0);
    int tc_ = 0;  // running tile counter -> TMEM slot
    for (int it = 0; it < my_items; ++it) {
      const int ntile = (item_of(it) & 1) ? 3 : 4;
      for (int kc = 0; kc < 2; ++kc) {
        mbar_wait(&pfull[kc], it & 1);
        tc_fence_after();
        const uint32_t pb = smem_u32(patch + kc * PATCH_BYTES);
        for (int t = 0; t < ntile; ++t) {
          const int slot = (tc_ + t) & 3;
          if (kc == 0) {
            mbar_wait(&tempty[slot], (((tc_ + t) >> 2) & 1) ^ 1);
            tc_fence_after();
          }
          const uint32_t d = tmem_base + slot * 128;
#pragma unroll
          for (int kh = 0; kh < 3; ++kh) {
            const uint32_t sa = pb + (uint32_t)((4 * t + 2 - kh) * ROW_BYTES);
            const uint32_t sw = smem_u32(wsm + kh * 3 * TAP_BYTES + kc * 4096);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              if (leader)
                umma_tf32(d, desc_kmajor(sa + kk * 32), desc_mnmajor(sw + kk * 1024, TAP_BYTES), idesc,
                          (kc | kh | kk) != 0);
          }
          if (kc == 1 && leader) umma_commit(&tfull[slot]);
        }
        if (leader) umma_commit(&pempty[kc]);
      }
      tc_ += ntile;
    }
  } else {
    // ---- 8 epilogue warps: warp pairs (w, w+4) share a TMEM sub-partition and split the 32 channels
    const int q = warp_id & 3, hc = warp_id >> 2;
    int tc_ = 0;
    for (int it = 0; it < my_items; ++it) {
      const int item = item_of(it), n = item >> 1, half = item & 1, y0 = half * 16;
      const int ntile = half ? 3 : 4;
      for (int t = 0; t < ntile; ++t) {
        const int slot = (tc_ + t) & 3;
        const int y = y0 + 4 * t + q, x = lane;
        const bool valid = y < H1 && x < W1;
        const size_t e = (((size_t)n * H1 + (valid ? y : 0)) * W1 + (valid ? x : 0)) * C1 + hc * 16;
        float4 am[4];  // ReLU mask operand, fetched before the accumulator is waited for
#pragma unroll
        for (int j = 0; j < 4; ++j) am[j] = *reinterpret_cast<const float4*>(a1 + e + 4 * j);
        mbar_wait(&tfull[slot], ((tc_ + t) >> 2) & 1);
        tc_fence_after();
        float v0[16], v1[16], v2[16];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + slot * 128 + hc * 16;
        tmem_ld16(taddr, v0);
        tmem_ld16(taddr + 32, v1);
        tmem_ld16(taddr + 64, v2);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[slot]);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float s1 = __shfl_up_sync(0xffffffffu, v1[j], 1);
          const float s2 = __shfl_up_sync(0xffffffffu, v2[j], 2);
          v0[j] += (lane >= 1 ? s1 : 0.f) + (lane >= 2 ? s2 : 0.f);
        }
        if (valid) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const float4 a = am[j >> 2];
            *reinterpret_cast<float4*>(da1 + e + j) =
                make_float4(a.x > 0.f ? v0[j] : 0.f, a.y > 0.f ? v0[j + 1] : 0.f, a.z > 0.f ? v0[j + 2] : 0.f,
                            a.w > 0.f ? v0[j + 3] : 0.f);
          }
        }
      }
      tc_ += ntile;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (is_mma) tmem_dealloc(tmem_base, 512);
}

// =============================================================================================
// forward v4 (DRAFT, same status as the dgrad above): conv1 computed inside conv2's producer warps.
// conv1 is 9 FMAs per output, so instead of a separate kernel writing a1 and TMA reading it back, four
// producer warps build the 26-row x 16-pixel x 32-channel a1 patch of an item directly in shared memory
// (same swizzled image as TMA would leave: offset = r*2048 + px*128 + ((cq ^ (px & 7)) << 4)) from the
// uint8 image, and also store it to global memory once for the backward pass.  MMA issue and epilogue are
// those of conv2_fwd_tc3_kernel<8>.
// =============================================================================================
constexpr int F_PITCH = 16, F_ROWS = 26;
constexpr int F_ROW_BYTES = F_PITCH * 128;
constexpr int F_PATCH_BYTES = F_ROWS * F_ROW_BYTES;
constexpr int F4_SMEM = 9 * TAP_BYTES + 2 * F_PATCH_BYTES + 1024;
constexpr int F4_THREADS = (EPI_WARPS + PROD_WARPS + 1) * 32;  // 416

__device__ __forceinline__ float normalize_px4(uint8_t v) {
  // same operation order as ToTensor (div 255) then Normalize (sub mean, div std) -- see conv1.cu
  return __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), MEAN), STD);
}

// 32 channels of one output pixel per lane -> 2x2 max-pool over (lane^1, lane^16) -> 8 channels of one
// pooled pixel per lane -> bias, ReLU, dropout, store (copy of tc_conv2_v3.cu: pool_store<32>)
__device__ __forceinline__ void pool32_store4(const float (&v)[32], int lane, int oc, const float* bias_s, int train,
                                              unsigned long long seed, unsigned long long step, uint32_t e_pixel,
                                              bool store, float* __restrict__ pooled, uint8_t* __restrict__ amax) {
  const bool b0 = lane & 1, b4 = (lane >> 4) & 1;
  float m1[16];
  uint32_t dxmask = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float mine = b0 ? v[16 + j] : v[j], send = b0 ? v[j] : v[16 + j];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
    const float ev = b0 ? recv : mine, od = b0 ? mine : recv;
    const bool gt = od > ev;  // strict: first maximum wins
    m1[j] = gt ? od : ev;
    dxmask |= (uint32_t)gt << j;
  }
  const uint32_t dx_other = __shfl_xor_sync(0xffffffffu, dxmask, 16);
  float m2[8];
  uint32_t am[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float mine = b4 ? m1[8 + j] : m1[j], send = b4 ? m1[j] : m1[8 + j];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 16);
    const int jj = (b4 ? 8 : 0) + j;
    const uint32_t dxm = (dxmask >> jj) & 1u, dxo = (dx_other >> jj) & 1u;
    const float top = b4 ? recv : mine, bot = b4 ? mine : recv;
    const uint32_t dxt = b4 ? dxo : dxm, dxb = b4 ? dxm : dxo;
    const bool gt = bot > top;
    m2[j] = gt ? bot : top;
    am[j] = gt ? (2u + dxb) : dxt;
  }
  if (!store) return;
  const int c0 = oc * 32 + (b0 ? 16 : 0) + (b4 ? 8 : 0);
  const uint32_t e = e_pixel + c0;
#pragma unroll
  for (int g4 = 0; g4 < 2; ++g4) {
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) o[c] = fmaxf(m2[g4 * 4 + c] + bias_s[c0 + g4 * 4 + c], 0.f);
    if (train & 2) {
      const uint4 r = dropout_bits4(seed, step, 1u, (e >> 2) + g4);
      const float sc = 1.0f / (1.0f - P_DROP1);
      o[0] = dropout_keep(r.x, DROP1_THRESH) ? o[0] * sc : 0.f;
      o[1] = dropout_keep(r.y, DROP1_THRESH) ? o[1] * sc : 0.f;
      o[2] = dropout_keep(r.z, DROP1_THRESH) ? o[2] * sc : 0.f;
      o[3] = dropout_keep(r.w, DROP1_THRESH) ? o[3] * sc : 0.f;
    }
    *reinterpret_cast<float4*>(pooled + e + g4 * 4) = make_float4(o[0], o[1], o[2], o[3]);
  }
  uint2 pk;
  pk.x = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
  pk.y = am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24);
  *reinterpret_cast<uint2*>(amax + e) = pk;
}

__global__ void __launch_bounds__(F4_THREADS, 1)
conv12_fwd_tc4_kernel(const __grid_constant__ CUtensorMap mapW, const uint8_t* __restrict__ images,
                      const int32_t* __restrict__ idx, const B200Ctrl* __restrict__ ctrl,
                      const float* __restrict__ w1, const float* __restrict__ b1v,  // conv1 [9][32], [32]
                      float* __restrict__ a1,                                        // [B][26][26][32] (for backward)
                      const float* __restrict__ bias, float* __restrict__ pooled, uint8_t* __restrict__ amax,
                      int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                        // [kh][kw][64 o][32 i]: per kh a K-major 192 x 32 B operand
  uint8_t* patch = smem + 9 * TAP_BYTES;      // [2][26 rows][16 px][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[2], tempty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[C2];
  __shared__ float xs[IMG * IMG];             // the item's normalised image (producer warps only)

  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_epi = warp_id < EPI_WARPS, is_mma = warp_id == EPI_WARPS + PROD_WARPS;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&pfull[i], PROD_WARPS); mbar_init(&pempty[i], 1);
      mbar_init(&tfull[i], 1); mbar_init(&tempty[i], EPI_WARPS);
    }
    fence_barrier_init();
  }
  if (is_mma) tmem_alloc(&tmem_base_smem, 512);  // 2 accumulator slots of 256 columns (192 used)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();
  if (threadIdx.x < C2) bias_s[threadIdx.x] = bias[threadIdx.x];
  __syncthreads();
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (!is_epi && !is_mma) {
    // ---- producers: conv2 weights by TMA (once); per item conv1 (+bias+ReLU) into the patch and into a1
    const int ptid = threadIdx.x - EPI_WARPS * 32;  // 0..127
    if (ptid < 32) {
      const bool leader = elect_one();
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t)
        if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
    }
    const int cq = ptid & 7;  // channel quad of this thread: fixed, so its conv1 weights live in registers
    float wr[9][4], br[4];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float4 v = *reinterpret_cast<const float4*>(w1 + t * C1 + cq * 4);
      wr[t][0] = v.x; wr[t][1] = v.y; wr[t][2] = v.z; wr[t][3] = v.w;
    }
    {
      const float4 v = *reinterpret_cast<const float4*>(b1v + cq * 4);
      br[0] = v.x; br[1] = v.y; br[2] = v.z; br[3] = v.w;
    }
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item >> 1, half = item & 1, X0 = half * 10, buf = it & 1;
      const int pos = ctrl->pos + n;
      const int sample = idx ? idx[pos] : pos;
      asm volatile("bar.sync 1, 128;" ::: "memory");  // everybody is done reading the previous image
      if (ptid < IMG * IMG / 4) {
        const uchar4 v = reinterpret_cast<const uchar4*>(images + (size_t)sample * (IMG * IMG))[ptid];
        xs[ptid * 4 + 0] = normalize_px4(v.x);
        xs[ptid * 4 + 1] = normalize_px4(v.y);
        xs[ptid * 4 + 2] = normalize_px4(v.z);
        xs[ptid * 4 + 3] = normalize_px4(v.w);
      }
      // 196 uchar4 for 128 threads: second round
      if (ptid + 128 < IMG * IMG / 4) {
        const int q = ptid + 128;
        const uchar4 v = reinterpret_cast<const uchar4*>(images + (size_t)sample * (IMG * IMG))[q];
        xs[q * 4 + 0] = normalize_px4(v.x);
        xs[q * 4 + 1] = normalize_px4(v.y);
        xs[q * 4 + 2] = normalize_px4(v.z);
        xs[q * 4 + 3] = normalize_px4(v.w);
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(&pempty[buf], ((it >> 1) & 1) ^ 1);
      uint8_t* pb = patch + buf * F_PATCH_BYTES;
      float* outg = a1 + (size_t)n * (H1 * W1 * C1);
      // work unit = (row r, pixel px, channel quad cq): 26 x 16 x 8; u & 7 == cq because the stride is 128
      for (int u = ptid; u < F_ROWS * F_PITCH * 8; u += PROD_WARPS * 32) {
        const int px = (u >> 3) & (F_PITCH - 1), r = u >> 7, col = X0 + px;
        float acc[4] = {br[0], br[1], br[2], br[3]};
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) {
            const float v = xs[(r + kh) * IMG + col + kw];
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[c] = fmaf(v, wr[kh * 3 + kw][c], acc[c]);
          }
        const float4 o = make_float4(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
        *reinterpret_cast<float4*>(pb + r * F_ROW_BYTES + px * 128 + ((cq ^ (px & 7)) << 4)) = o;
        // the halves overlap in columns 10..15: half 0 owns columns 0..12, half 1 columns 13..25
        if (half ? (px >= 3) : (px <= 12)) *reinterpret_cast<float4*>(outg + ((size_t)r * W1 + col) * C1 + cq * 4) = o;
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&pfull[buf]);
    }
  } else if (is_mma) {
    const bool leader = elect_one();
    constexpr uint32_t idesc = idesc_tf32(128, 192, 0, 0);
    mbar_wait(&w_bar, 0);
    int tc_ = 0;
    for (int it = 0; it < my_items; ++it) {
      const int buf = it & 1;
      mbar_wait(&pfull[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t pb = smem_u32(patch + buf * F_PATCH_BYTES);
      for (int ty = 0; ty < 3; ++ty, ++tc_) {
        const int slot = tc_ & 1;
        mbar_wait(&tempty[slot], ((tc_ >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + slot * 256;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
          const uint32_t sa = pb + (uint32_t)((ty * 8 + kh) * F_ROW_BYTES);
          const uint32_t sw = smem_u32(wsm + kh * 3 * TAP_BYTES);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            if (leader) umma_tf32(d, desc_kmajor(sa + kk * 32), desc_kmajor(sw + kk * 32), idesc, (kh | kk) != 0);
        }
        if (leader) umma_commit(&tfull[slot]);
      }
      if (leader) umma_commit(&pempty[buf]);
    }
  } else {
    // ---- 8 epilogue warps: warp pairs (w, w+4) share a TMEM sub-partition; each takes 32 of the 64 channels
    const int q = warp_id & 3, oc = warp_id >> 2;
    const int train = ctrl->train;
    const unsigned long long seed = ctrl->seed, step = ctrl->step;
    const int lx = lane & 15;
    int tc_ = 0;
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item >> 1, half = item & 1, X0 = half * 10;
      const bool store = half ? (lx >= 2 && lx < 14) : (lx < 12);
      for (int ty = 0; ty < 3; ++ty, ++tc_) {
        const int slot = tc_ & 1;
        mbar_wait(&tfull[slot], (tc_ >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + slot * 256;
        const int ph = ty * 4 + q, pw = (X0 + lx) >> 1;
        const uint32_t e_pixel = ((uint32_t)(n * HP + ph) * WP + pw) * C2;
        float v0[32], v1[32], v2[32];
        tmem_ld32(taddr + oc * 32, v0);        // kw = 0
        tmem_ld32(taddr + 64 + oc * 32, v1);   // kw = 1
        tmem_ld32(taddr + 128 + oc * 32, v2);  // kw = 2
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[slot]);
#pragma unroll
        for (int j = 0; j < 32; ++j)
          v0[j] += __shfl_down_sync(0xffffffffu, v1[j], 1) + __shfl_down_sync(0xffffffffu, v2[j], 2);
        pool32_store4(v0, lane, oc, bias_s, train, seed, step, e_pixel, store, pooled, amax);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (is_mma) tmem_dealloc(tmem_base, 512);
}

std::map<const void*, CUtensorMap> g_wmaps;
std::mutex g_mu;

// atom32 = 1: 32B-atom swizzle (N-major B operand of the dgrad); 0: plain 128B swizzle (K-major, forward)
int weight_map(const float* w, int atom32, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  const void* key = (const char*)w + atom32;  // weights are 256-float aligned: the low bit is free
  auto it = g_wmaps.find(key);
  if (it != g_wmaps.end()) { *out = it->second; return B200_OK; }
  CUtensorMap m;
  uint64_t dims[2] = {32, 576}, strides[1] = {128};
  uint32_t box[2] = {32, 64};
  int r = tcmap::encode(&m, w, 2, dims, strides, box, atom32);
  if (r != B200_OK) return r;
  g_wmaps[key] = m;
  *out = m;
  return B200_OK;
}

}  // namespace

// DRAFT entry point (see the file header): da1 = relu'(a1) * conv2_dgrad(max-pool-backward(dp, amax))
extern "C" int b200_conv2_dgrad_tc4(const float* dp, const uint8_t* amax, const float* w, const float* a1, float* da1,
                                    int batch, void* stream) {
  B200_REQUIRE(batch > 0 && dp && amax && w && a1 && da1, "conv2_dgrad_tc4: bad arguments");
  CUtensorMap mw;
  int r = weight_map(w, 1, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_tc4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DG4_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  B200_CHECK(b200_launch(conv2_dgrad_tc4_kernel, dim3(grid), dim3(DG4_THREADS), (size_t)DG4_SMEM, (cudaStream_t)stream, mw,
                         dp, amax, a1, da1, nitems));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// DRAFT entry point: conv1 (gather + normalise + conv + bias + ReLU) and conv2 (+ bias + ReLU + 2x2 max-pool +
// dropout) in one kernel; a1 is still written once because the backward pass reads it.
extern "C" int b200_conv12_fwd_tc4(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* w1,
                                   const float* b1, float* a1, const float* w2, const float* b2, float* pooled,
                                   uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0 && images && ctrl && w1 && b1 && a1 && w2 && b2 && pooled && amax, "conv12_fwd_tc4: bad arguments");
  CUtensorMap mw;
  int r = weight_map(w2, 0, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv12_fwd_tc4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F4_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  B200_CHECK(b200_launch_serial(conv12_fwd_tc4_kernel, dim3(grid), dim3(F4_THREADS), (size_t)F4_SMEM, (cudaStream_t)stream, mw,
                                images, idx, (const B200Ctrl*)ctrl, w1, b1, a1, b2, pooled, amax, nitems));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
