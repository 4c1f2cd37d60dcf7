// conv2 forward on the 5th-generation tensor cores (mnist_ddp.py:50-55).
//
// "kw-scatter" implicit GEMM.  Measured on B200 (profiles/README.md): an SS-mode 128 x N x 8 tf32 tcgen05.mma
// costs about (A bytes + B bytes) / 64 cycles -- the operand read from shared memory, not the tensor pipe,
// sets the pace when N is small.  So the three kw taps sit on the N side of the MMA:
//
//   D[(y,x')][(kw,o)] = sum_kh sum_i a1[y+kh][x'][i] * W[kh][kw][o][i]      (N = 192)
//   out[y][x][o]      = sum_kw D[(y, x+kw)][(kw,o)]   -> bias, ReLU, 2x2 max-pool, dropout
//
// Item = (image, x-half): a 26-row x 16-pixel patch of a1 (columns 0..15 or 10..25, 52 KB, pixel = one
// 128-byte swizzle row) feeds three 8-row x 16-pixel tiles; a tile is a contiguous 16 KiB K-major operand
// at patch row (8*ty + kh): a 3x3 tap is an address offset, not a new load, because the tensor core
// applies the 128B swizzle on absolute shared-memory address bits exactly like TMA does when it writes
// (DESIGN.md section 3).  The B operand of one kh is the native weight block [kw][o][i] = 192 rows of
// 128 bytes.  12 MMAs of 128x192x8 per tile.  A warp holds two tile rows (lanes 0..15 = row 2q, lanes
// 16..31 = row 2q+1), so the kw-sum is a shuffle-down by 1 and 2 and the 2x2 pool is the (lane^1, lane^16)
// neighbourhood.  14 of the 16 columns are valid outputs; the two halves overlap.
//
// Two kernels share the MMA issue loop and the epilogue:
//   conv2_fwd_tc_kernel   : a1 comes from memory, patch rows loaded by TMA        (b200_conv2_fwd_tc)
//   conv12_fwd_tc_kernel  : conv1 is computed INSIDE the producer warps -- 9 FMAs per output from the uint8
//                           image (index gather + ToTensor/Normalize + conv + bias + ReLU), written both into
//                           the swizzled patch and once to memory for the backward pass.  This is the
//                           engine's forward: one launch from the resident dataset to the pooled activation.
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
constexpr int F_PITCH = 16, F_ROWS = 26;
constexpr int F_ROW_BYTES = F_PITCH * 128;             // 2 KiB
constexpr int F_PATCH_BYTES = F_ROWS * F_ROW_BYTES;    // 53,248
constexpr int F3_SMEM = 9 * TAP_BYTES + 2 * F_PATCH_BYTES + 1024;

// NCH (32 or 16) channels of one output pixel per lane -> 2x2 max-pool over (lane^1, lane^16) -> NCH/4
// channels of one pooled pixel per lane (channel offset c0) -> bias, ReLU, dropout, store.
template <int NCH>
__device__ __forceinline__ void pool_store(const float (&v)[NCH], int lane, int oc, const float* bias_s, int train,
                                           unsigned long long seed, unsigned long long step, uint32_t e_pixel,
                                           bool store, float* __restrict__ pooled, uint8_t* __restrict__ amax) {
  constexpr int HA = NCH / 2, HB = NCH / 4;
  const bool b0 = lane & 1, b4 = (lane >> 4) & 1;
  float m1[HA];
  uint32_t dxmask = 0;
#pragma unroll
  for (int j = 0; j < HA; ++j) {
    const float mine = b0 ? v[HA + j] : v[j], send = b0 ? v[j] : v[HA + j];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
    const float ev = b0 ? recv : mine, od = b0 ? mine : recv;
    const bool gt = od > ev;  // strict: first maximum wins
    m1[j] = gt ? od : ev;
    dxmask |= (uint32_t)gt << j;
  }
  const uint32_t dx_other = __shfl_xor_sync(0xffffffffu, dxmask, 16);
  float m2[HB];
  uint32_t am[HB];
#pragma unroll
  for (int j = 0; j < HB; ++j) {
    const float mine = b4 ? m1[HB + j] : m1[j], send = b4 ? m1[j] : m1[HB + j];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 16);
    const int jj = (b4 ? HB : 0) + j;
    const uint32_t dxm = (dxmask >> jj) & 1u, dxo = (dx_other >> jj) & 1u;
    const float top = b4 ? recv : mine, bot = b4 ? mine : recv;
    const uint32_t dxt = b4 ? dxo : dxm, dxb = b4 ? dxm : dxo;
    const bool gt = bot > top;
    m2[j] = gt ? bot : top;
    am[j] = gt ? (2u + dxb) : dxt;
  }
  if (!store) return;
  const int c0 = oc * NCH + (b0 ? HA : 0) + (b4 ? HB : 0);
  const uint32_t e = e_pixel + c0;
#pragma unroll
  for (int g4 = 0; g4 < HB / 4; ++g4) {
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
  if (HB == 8) {
    uint2 pk;
    pk.x = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
    pk.y = am[HB - 4] | (am[HB - 3] << 8) | (am[HB - 2] << 16) | (am[HB - 1] << 24);
    *reinterpret_cast<uint2*>(amax + e) = pk;
  } else {
    *reinterpret_cast<uint32_t*>(amax + e) = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
  }
}

// EW = number of epilogue warps: 8 (32 channels per warp) or 16 (16 channels per warp)
template <int EW>
__global__ void __launch_bounds__(EW * 32 + 64, 1)
conv2_fwd_tc_kernel(const __grid_constant__ CUtensorMap mapP, const __grid_constant__ CUtensorMap mapW,
                     const float* __restrict__ bias, const B200Ctrl* __restrict__ ctrl, float* __restrict__ pooled,
                     uint8_t* __restrict__ amax, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                        // [kh][kw][64 o][32 i]: per kh a K-major 192 x 32 B operand
  uint8_t* patch = smem + 9 * TAP_BYTES;      // [2][26 rows][16 px][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[2], tempty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[C2];

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapP);
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&pfull[i], 1); mbar_init(&pempty[i], 1);
      mbar_init(&tfull[i], 1); mbar_init(&tempty[i], EW);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, 512);  // 2 accumulator slots of 256 columns (192 used)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();  // on-chip set-up above overlaps the previous kernel's tail
  if (threadIdx.x < C2) bias_s[threadIdx.x] = bias[threadIdx.x];
  __syncthreads();
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    const bool leader = elect_one();
    if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
    for (int t = 0; t < 9; ++t)
      if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item >> 1, X0 = (item & 1) * 10, buf = it & 1;
      mbar_wait(&pempty[buf], ((it >> 1) & 1) ^ 1);
      if (leader) mbar_expect_tx(&pfull[buf], F_PATCH_BYTES);
      for (int r = 0; r < F_ROWS; ++r)
        if (leader) tma_load_4d(patch + buf * F_PATCH_BYTES + r * F_ROW_BYTES, &mapP, &pfull[buf], 0, X0, r, n);
    }
  } else if (warp == 1) {
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
    // warps w, w+4, .. share a TMEM sub-partition; each takes NCH of the 64 channels
    constexpr int NCH = 256 / EW;
    const int q = wq, oc = (warp - 2) >> 2;
    const int train = ctrl->train;
    const unsigned long long seed = ctrl->seed, step = ctrl->step;
    const int lx = lane & 15;
    int tc_ = 0;
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item >> 1, half = item & 1, X0 = half * 10;
      // half 0 stores output columns 0..11 (local 0..11), half 1 columns 12..23 (local 2..13)
      const bool store = half ? (lx >= 2 && lx < 14) : (lx < 12);
      for (int ty = 0; ty < 3; ++ty, ++tc_) {
        const int slot = tc_ & 1;
        mbar_wait(&tfull[slot], (tc_ >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + slot * 256;
        const int ph = ty * 4 + q, pw = (X0 + lx) >> 1;
        const uint32_t e_pixel = ((uint32_t)(n * HP + ph) * WP + pw) * C2;
        float v0[NCH], v1[NCH], v2[NCH];
        if (NCH == 32) {
          tmem_ld32(taddr + oc * 32, reinterpret_cast<float(&)[32]>(v0));        // kw = 0
          tmem_ld32(taddr + 64 + oc * 32, reinterpret_cast<float(&)[32]>(v1));   // kw = 1
          tmem_ld32(taddr + 128 + oc * 32, reinterpret_cast<float(&)[32]>(v2));  // kw = 2
        } else {
          tmem_ld16(taddr + oc * 16, reinterpret_cast<float(&)[16]>(v0));
          tmem_ld16(taddr + 64 + oc * 16, reinterpret_cast<float(&)[16]>(v1));
          tmem_ld16(taddr + 128 + oc * 16, reinterpret_cast<float(&)[16]>(v2));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[slot]);
#pragma unroll
        for (int j = 0; j < NCH; ++j)
          v0[j] += __shfl_down_sync(0xffffffffu, v1[j], 1) + __shfl_down_sync(0xffffffffu, v2[j], 2);
        pool_store<NCH>(v0, lane, oc, bias_s, train, seed, step, e_pixel, store, pooled, amax);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// =============================================================================================
// conv1 inside conv2's producer warps.  Four producer warps build the 26-row x 16-pixel x 32-channel a1 patch
// of an item directly in shared memory -- the same swizzled image TMA would leave:
//     offset = r*2048 + px*128 + ((cq ^ (px & 7)) << 4)        (cq = channel quad)
// -- from the normalised uint8 image, make it visible to the tensor core with fence.proxy.async, and also
// store it to global memory once for the backward pass.  FMA order is that of conv1_fwd_kernel (conv1.cu), so
// a1 is bit-identical to the two-kernel path.  MMA issue and epilogue are those of conv2_fwd_tc_kernel<8>.
// =============================================================================================
constexpr int F4_SMEM = 9 * TAP_BYTES + 2 * F_PATCH_BYTES + 1024;
constexpr int EPI_WARPS = 8, PROD_WARPS = 4;
constexpr int F4_THREADS = (EPI_WARPS + PROD_WARPS + 1) * 32;  // 416

__device__ __forceinline__ float normalize_px4(uint8_t v) {
  // same operation order as ToTensor (div 255) then Normalize (sub mean, div std) -- see conv1.cu
  return __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), MEAN), STD);
}

__global__ void __launch_bounds__(F4_THREADS, 1)
conv12_fwd_tc_kernel(const __grid_constant__ CUtensorMap mapW, const uint8_t* __restrict__ images,
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
        pool_store<32>(v0, lane, oc, bias_s, train, seed, step, e_pixel, store, pooled, amax);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (is_mma) tmem_dealloc(tmem_base, 512);
}

// tensor maps, cached per (pointer, batch): kind 0 = a1 patch row (32 channels x 16 px x 1 row), 1 = weights (K-major)
struct Key {
  const void* p; int batch, kind;
  bool operator<(const Key& o) const { return std::tie(p, batch, kind) < std::tie(o.p, o.batch, o.kind); }
};
std::map<Key, CUtensorMap> g_maps;
std::mutex g_mu;

int get_map(const float* p, int batch, int kind, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  Key key{p, batch, kind};
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return B200_OK; }
  if (g_maps.size() >= 64) g_maps.clear();  // bounded: a long-lived host may cycle through many buffers
  CUtensorMap m;
  int r;
  if (kind == 0) {
    uint64_t dims[4] = {32, 26, 26, (uint64_t)batch}, strides[3] = {128, 26 * 128, 676 * 128};
    uint32_t box[4] = {32, F_PITCH, 1, 1};
    r = tcmap::encode(&m, p, 4, dims, strides, box, 0);
  } else {
    uint64_t dims[2] = {32, 576}, strides[1] = {128};
    uint32_t box[2] = {32, 64};
    r = tcmap::encode(&m, p, 2, dims, strides, box, 0);
  }
  if (r != B200_OK) return r;
  g_maps[key] = m;
  *out = m;
  return B200_OK;
}

}  // namespace

/* implicit-GEMM conv2 forward with the fused bias+ReLU+pool+dropout epilogue (same outputs as _simt) */
extern "C" int b200_conv2_fwd_tc(const float* a1, const float* w, const float* bias, const void* ctrl, float* pooled,
                                 uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0 && a1 && w && bias && ctrl && pooled && amax, "conv2_fwd_tc: bad arguments");
  CUtensorMap mp, mw;
  int r = get_map(a1, batch, 0, &mp);
  if (r != B200_OK) return r;
  r = get_map(w, 0, 1, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_tc_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_tc_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  static const int ew = getenv("B200_FWD_EW") ? atoi(getenv("B200_FWD_EW")) : 16;
  if (ew == 8)
    B200_CHECK(b200_launch(conv2_fwd_tc_kernel<8>, dim3(grid), dim3(320), (size_t)F3_SMEM, (cudaStream_t)stream, mp, mw, bias,
                           (const B200Ctrl*)ctrl, pooled, amax, nitems));
  else
    B200_CHECK(b200_launch(conv2_fwd_tc_kernel<16>, dim3(grid), dim3(576), (size_t)F3_SMEM, (cudaStream_t)stream, mp, mw, bias,
                           (const B200Ctrl*)ctrl, pooled, amax, nitems));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// conv1 (gather + normalise + conv + bias + ReLU) and conv2 (+ bias + ReLU + 2x2 max-pool + dropout) in one
// kernel; a1 is still written once because the backward pass reads it.
extern "C" int b200_conv12_fwd_tc(const uint8_t* images, const int32_t* idx, const void* ctrl, const float* w1,
                                  const float* b1, float* a1, const float* w2, const float* b2, float* pooled,
                                  uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0 && images && ctrl && w1 && b1 && a1 && w2 && b2 && pooled && amax, "conv12_fwd_tc: bad arguments");
  CUtensorMap mw;
  int r = get_map(w2, 0, 1, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv12_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F4_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  B200_CHECK(b200_launch_serial(conv12_fwd_tc_kernel, dim3(grid), dim3(F4_THREADS), (size_t)F4_SMEM, (cudaStream_t)stream, mw,
                                images, idx, (const B200Ctrl*)ctrl, w1, b1, a1, b2, pooled, amax, nitems));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
