// conv2 forward, second generation: ONE TMA load per (image, half) instead of one per tap.
//
// tc_conv2.cu's forward re-reads every input pixel nine times from L2 (one shifted TMA box per tap)
// and is L2-bandwidth bound (ncu: ~28 us at batch 200 against ~5 us of tensor-core work).  Here the
// producer loads an 18-row x 32-pixel x 32-channel input patch once (73.7 KB, pixel = 128-byte
// swizzle row, row pitch padded to 32 pixels so that every output row starts on a 4 KiB boundary).
// A 3x3 tap is then not a new load but a new VIEW of the same patch: the A-operand descriptor starts
// (kh*32 + x0 + kw) rows into the patch, its 16 eight-row groups (16 output rows x 8 output pixels)
// are 4096 bytes apart (the tensor core swizzles on absolute smem address bits, as TMA does, so a
// row-shifted start needs no further adjustment).  Per (image, half): 3 x-blocks x 9 taps x 4 k-steps = 108 tcgen05.mma of
// 128x64x8, three TMEM accumulators (double buffered across items), same fused epilogue as v1
// (2x2 max-pool by warp shuffles, bias, ReLU, Philox dropout, 2-bit argmax).
//   half 0 computes output rows 0..15, half 1 rows 8..23 and stores rows 16..23 only, so no patch
//   ever reaches outside the 26-row input.
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
constexpr int ROW_BYTES = P_PITCH * 128;             // one patch row = one TMA request (4 KiB)
constexpr int PATCH_BYTES = P_ROWS * ROW_BYTES;      // 73,728
constexpr int V2_SMEM = 9 * TAP_BYTES + 2 * PATCH_BYTES + 1024;

// 2x2 max-pool over the (lane^1, lane^8) neighbourhood of a 128x64 accumulator tile held as
// v0 = columns 0..31, v1 = columns 32..63 of this lane's row, then bias + ReLU + dropout + store.
// After the two exchanges a lane owns 16 channels (c0..c0+15) of one pooled pixel.
__device__ __forceinline__ void pool_bias_relu_dropout_store(const float (&v0)[32], const float (&v1)[32], int lane,
                                                             const float* bias_s, int train, unsigned long long seed,
                                                             unsigned long long step, uint32_t e_pixel, bool store,
                                                             float* __restrict__ pooled, uint8_t* __restrict__ amax) {
  const bool b0 = lane & 1, b3 = (lane >> 3) & 1;
  const int c0 = (b0 ? 32 : 0) + (b3 ? 16 : 0);
  float m1[32];
  uint32_t dxmask = 0;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float mine = b0 ? v1[j] : v0[j], send = b0 ? v0[j] : v1[j];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
    const float ev = b0 ? recv : mine, od = b0 ? mine : recv;
    const bool gt = od > ev;  // strict: the first maximum wins, like ATen's max_pool2d
    m1[j] = gt ? od : ev;
    dxmask |= (uint32_t)gt << j;
  }
  const uint32_t dx_other = __shfl_xor_sync(0xffffffffu, dxmask, 8);
  float m2[16];
  uint32_t am[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float mine = b3 ? m1[16 + j] : m1[j], send = b3 ? m1[j] : m1[16 + j];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 8);
    const int jj = (b3 ? 16 : 0) + j;
    const uint32_t dxm = (dxmask >> jj) & 1u, dxo = (dx_other >> jj) & 1u;
    const float top = b3 ? recv : mine, bot = b3 ? mine : recv;
    const uint32_t dxt = b3 ? dxo : dxm, dxb = b3 ? dxm : dxo;
    const bool gt = bot > top;
    m2[j] = gt ? bot : top;
    am[j] = gt ? (2u + dxb) : dxt;
  }
  if (!store) return;
  const uint32_t e = e_pixel + c0;
#pragma unroll
  for (int g4 = 0; g4 < 4; ++g4) {
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
  uint4 pk;
  pk.x = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
  pk.y = am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24);
  pk.z = am[8] | (am[9] << 8) | (am[10] << 16) | (am[11] << 24);
  pk.w = am[12] | (am[13] << 8) | (am[14] << 16) | (am[15] << 24);
  *reinterpret_cast<uint4*>(amax + e) = pk;
}

__global__ void __launch_bounds__(192, 1)
conv2_fwd_tc2_kernel(const __grid_constant__ CUtensorMap mapP, const __grid_constant__ CUtensorMap mapW,
                     const float* __restrict__ bias, const B200Ctrl* __restrict__ ctrl, float* __restrict__ pooled,
                     uint8_t* __restrict__ amax, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                        // [9][64 o][32 i]
  uint8_t* patch = smem + 9 * TAP_BYTES;      // [2][18 rows][32 px][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[6], tempty[6];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[C2];

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapP);
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&pfull[i], 1); mbar_init(&pempty[i], 1); }
    for (int i = 0; i < 6; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
  }
  if (threadIdx.x < C2) bias_s[threadIdx.x] = bias[threadIdx.x];
  if (warp == 1) tmem_alloc(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t) if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
      for (int it = 0; it < my_items; ++it) {
        const int item = blockIdx.x + it * gridDim.x, n = item >> 1, y0 = (item & 1) * 8, buf = it & 1;
        mbar_wait(&pempty[buf], ((it >> 1) & 1) ^ 1);
        if (leader) mbar_expect_tx(&pfull[buf], PATCH_BYTES);
        for (int r = 0; r < P_ROWS; ++r)  // one request per patch row: many small TMA requests in flight
          if (leader) tma_load_4d(patch + buf * PATCH_BYTES + r * ROW_BYTES, &mapP, &pfull[buf], 0, 0, y0 + r, n);
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(128, 64, 0, 0);
      mbar_wait(&w_bar, 0);
      for (int it = 0; it < my_items; ++it) {
        const int buf = it & 1;
        mbar_wait(&pfull[buf], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t pb = smem_u32(patch + buf * PATCH_BYTES);
        for (int xb = 0; xb < 3; ++xb) {
          const int slot = buf * 3 + xb;
          mbar_wait(&tempty[slot], ((it >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d = tmem_base + slot * 64;
#pragma unroll 1
          for (int tap = 0; tap < 9; ++tap) {
            const int kh = tap / 3, kw = tap - kh * 3;
            const uint32_t sa = pb + (uint32_t)((kh * P_PITCH + xb * 8 + kw) * 128);
            const uint32_t sw = smem_u32(wsm + tap * TAP_BYTES);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              if (leader) umma_tf32(d, desc_kmajor_view(sa + kk * 32, P_PITCH * 128), desc_kmajor(sw + kk * 32), idesc,
                        (tap | kk) != 0);
          }
          if (leader) umma_commit(&tfull[slot]);
        }
        if (leader) umma_commit(&pempty[buf]);  // every MMA reading this patch has completed
      }
    }
  } else {
    const int q = wq;
    const int train = ctrl->train;
    const unsigned long long seed = ctrl->seed, step = ctrl->step;
    const int py = q * 4 + (lane >> 3), px = lane & 7;
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item >> 1, half = item & 1, y0 = half * 8, buf = it & 1;
      const bool mine = (half == 0) || (py >= 8);  // half 1 recomputes rows 8..15: not stored again
      for (int xb = 0; xb < 3; ++xb) {
        const int slot = buf * 3 + xb;
        mbar_wait(&tfull[slot], (it >> 1) & 1);
        tc_fence_after();
        float v0[32], v1[32];
        const bool warp_stores = (half == 0) || (q >= 2);
        if (warp_stores) {
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + slot * 64;
          tmem_ld32(taddr, v0);
          tmem_ld32(taddr + 32, v1);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[slot]);
        if (warp_stores) {
          const int ph = (y0 + py) >> 1, pw = (xb * 8 + px) >> 1;
          const uint32_t e_pixel = ((uint32_t)(n * HP + ph) * WP + pw) * C2;
          pool_bias_relu_dropout_store(v0, v1, lane, bias_s, train, seed, step, e_pixel, mine, pooled, amax);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// =============================================================================================
// dgrad v2:  da1[y][x][i] = relu'(a1) * sum_{kh,kw,o} dY[y-kh][x-kw][o] * W[kh][kw][o][i]
// Item = (image, half): output rows y0..y0+15 (y0 = 0 or 10; half 1 stores rows 16..25 only).
// The 64 dY channels are two K-chunks of 32 (one 128-byte swizzle row each), loaded as two 18-row x
// 32-pixel patches starting at dY (row y0-2, column -2); everything outside dY is zero-filled by TMA.
// The two chunks form a 2-deep pipeline over (item, chunk): while the tensor core works on one chunk
// the producer already loads the next.  4 x-blocks (x0 = 0, 8, 16, 18) x 9 taps x 4 k-steps per chunk,
// 128x32x8 MMAs, weights are the N-major B operand ([tap][o][i], i contiguous).
// =============================================================================================
constexpr int DG_SMEM = 9 * TAP_BYTES + 2 * PATCH_BYTES + 1024;

__global__ void __launch_bounds__(192, 1)
conv2_dgrad_tc2_kernel(const __grid_constant__ CUtensorMap mapDY, const __grid_constant__ CUtensorMap mapW,
                       const float* __restrict__ a1, float* __restrict__ da1, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                     // [9][64 o][32 i], 32B-atom swizzle (N-major operand)
  uint8_t* patch = smem + 9 * TAP_BYTES;   // [2 stages][18][32][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[8], tempty[8];
  __shared__ uint32_t tmem_base_smem;

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapDY);
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&pfull[i], 1); mbar_init(&pempty[i], 1); }
    for (int i = 0; i < 8; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t) if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
      for (int it = 0; it < my_items; ++it) {
        const int item = blockIdx.x + it * gridDim.x, n = item >> 1, y0 = (item & 1) * 10;
        for (int kc = 0; kc < 2; ++kc) {  // stage == kc: chunk kc of every item goes through stage kc
          mbar_wait(&pempty[kc], (it & 1) ^ 1);
          if (leader) mbar_expect_tx(&pfull[kc], PATCH_BYTES);
          for (int r = 0; r < P_ROWS; ++r)
            if (leader) tma_load_4d(patch + kc * PATCH_BYTES + r * ROW_BYTES, &mapDY, &pfull[kc], kc * 32, -2, y0 - 2 + r, n);
        }
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(128, 32, 0, 1);
      mbar_wait(&w_bar, 0);
      for (int it = 0; it < my_items; ++it) {
        const int abuf = it & 1;
        for (int kc = 0; kc < 2; ++kc) {
          mbar_wait(&pfull[kc], it & 1);
          tc_fence_after();
          const uint32_t pb = smem_u32(patch + kc * PATCH_BYTES);
          for (int xb = 0; xb < 4; ++xb) {
            const int slot = abuf * 4 + xb, x0 = xb < 3 ? xb * 8 : 18;
            if (kc == 0) {
              mbar_wait(&tempty[slot], ((it >> 1) & 1) ^ 1);
              tc_fence_after();
            }
            const uint32_t d = tmem_base + slot * 32;
#pragma unroll 1
            for (int tap = 0; tap < 9; ++tap) {
              const int kh = tap / 3, kw = tap - kh * 3;
              const uint32_t sa = pb + (uint32_t)(((2 - kh) * P_PITCH + x0 + 2 - kw) * 128);
              const uint32_t sw = smem_u32(wsm + tap * TAP_BYTES + kc * 4096);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                if (leader) umma_tf32(d, desc_kmajor_view(sa + kk * 32, P_PITCH * 128), desc_mnmajor(sw + kk * 1024, TAP_BYTES),
                          idesc, (kc | tap | kk) != 0);
            }
            if (kc == 1) if (leader) umma_commit(&tfull[slot]);
          }
          if (leader) umma_commit(&pempty[kc]);
        }
      }
    }
  } else {
    const int q = wq;
    const int g = q * 4 + (lane >> 3), xx = lane & 7;
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item >> 1, half = item & 1, y0 = half * 10, abuf = it & 1;
      const int y = y0 + g;
      for (int xb = 0; xb < 4; ++xb) {
        const int slot = abuf * 4 + xb, x = (xb < 3 ? xb * 8 : 18) + xx;
        mbar_wait(&tfull[slot], (it >> 1) & 1);
        tc_fence_after();
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + slot * 32, v);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[slot]);
        const bool mine = (half == 0 || y >= 16) && (xb < 3 || x >= 24);  // skip the recomputed overlap
        if (mine) {
          const size_t e = (((size_t)n * H1 + y) * W1 + x) * C1;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 a = *reinterpret_cast<const float4*>(a1 + e + j);
            *reinterpret_cast<float4*>(da1 + e + j) =
                make_float4(a.x > 0.f ? v[j] : 0.f, a.y > 0.f ? v[j + 1] : 0.f, a.z > 0.f ? v[j + 2] : 0.f,
                            a.w > 0.f ? v[j + 3] : 0.f);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// =============================================================================================
// wgrad v2:  dW[kh][kw][o][i] = sum_pixels dY[y][x][o] * a1[y+kh][x+kw][i]      (K = pixels)
// Item = (image, third): 8 dY rows.  Loads: a1 rows y0..y0+9 (10 x 32 px patch) and the two 32-channel
// chunks of dY (8 x 32 px patches), all with the 32B-atom swizzle that MN-major tf32 operands need.
// One k-step = 8 consecutive pixels of one row.  A operand: M = (kw, i), four 32-wide chunks that are
// ONE PIXEL (128 bytes) apart in the a1 patch (kw = 3 is a don't-care chunk); B operand: N = 64
// channels, two chunks one patch apart.  Three accumulators (kh = 0..2) of 128 x 64 live in TMEM for the
// whole kernel; a CTA writes one partial at the end.
// =============================================================================================
constexpr int WG_A_BYTES = 10 * P_PITCH * 128;  // 40,960
constexpr int WG_Y_BYTES = 8 * P_PITCH * 128;   // 32,768 per channel chunk
constexpr int WG_STAGE = WG_A_BYTES + 2 * WG_Y_BYTES;  // 106,496
constexpr int WG_SMEM = 2 * WG_STAGE + 1024;
constexpr int CONV2_PART = 9 * C2 * C1 + C2;

__global__ void __launch_bounds__(192, 1)
conv2_wgrad_tc2_kernel(const __grid_constant__ CUtensorMap mapA1, const __grid_constant__ CUtensorMap mapDY,
                       float* __restrict__ part, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2], tfull_bar;
  __shared__ uint32_t tmem_base_smem;

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA1);
    tma_prefetch_desc(&mapDY);
    for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      for (int it = 0; it < my_items; ++it) {
        const int item = blockIdx.x + it * gridDim.x, n = item / 3, y0 = (item - n * 3) * 8, s = it & 1;
        mbar_wait(&empty_bar[s], ((it >> 1) & 1) ^ 1);
        uint8_t* st = smem + s * WG_STAGE;
        if (leader) mbar_expect_tx(&full_bar[s], WG_STAGE);
        for (int r = 0; r < 10; ++r) if (leader) tma_load_4d(st + r * ROW_BYTES, &mapA1, &full_bar[s], 0, 0, y0 + r, n);
        for (int r = 0; r < 8; ++r) {
          if (leader) tma_load_4d(st + WG_A_BYTES + r * ROW_BYTES, &mapDY, &full_bar[s], 0, 0, y0 + r, n);
          if (leader) tma_load_4d(st + WG_A_BYTES + WG_Y_BYTES + r * ROW_BYTES, &mapDY, &full_bar[s], 32, 0, y0 + r, n);
        }
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(128, 64, 1, 1);
      for (int it = 0; it < my_items; ++it) {
        const int s = it & 1;
        mbar_wait(&full_bar[s], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * WG_STAGE), sy = sa + WG_A_BYTES;
#pragma unroll 1
        for (int ks = 0; ks < 24; ++ks) {  // 8 rows x 3 eight-pixel blocks
          const int yy = ks / 3, x0 = (ks - yy * 3) * 8;
          const uint64_t bd = desc_mnmajor_view(sy + (uint32_t)((yy * P_PITCH + x0) * 128), WG_Y_BYTES, 512);
#pragma unroll
          for (int kh = 0; kh < 3; ++kh)
            if (leader) umma_tf32(tmem_base + kh * 64,
                      desc_mnmajor_view(sa + (uint32_t)(((yy + kh) * P_PITCH + x0) * 128), 128, 512), bd, idesc,
                      (it | ks) != 0);
        }
        if (leader) umma_commit(&empty_bar[s]);
      }
      if (leader) umma_commit(&tfull_bar);
    }
  } else {
    const int q = wq;  // = kw of the rows this warp reads (q == 3: don't-care chunk)
    mbar_wait(&tfull_bar, 0);
    tc_fence_after();
    float* dst = part + (size_t)blockIdx.x * CONV2_PART;
    for (int kh = 0; kh < 3; ++kh) {
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + kh * 64 + c0, v);
        if (q < 3) {
          const int tap = kh * 3 + q;
#pragma unroll
          for (int j = 0; j < 32; ++j) dst[(tap * C2 + c0 + j) * C1 + lane] = my_items ? v[j] : 0.f;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

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
  CUtensorMap m;
  int r;
  // kind: 0 a1 patch 18 rows (K-major)      1 weights (K-major)          2 dY patch 18 rows (K-major)
  //       3 weights (32B-atom, N-major)     4 a1 patch 10 rows (32B-atom) 5 dY patch 8 rows (32B-atom)
  if (kind == 1 || kind == 3) {
    uint64_t dims[2] = {32, 576}, strides[1] = {128};
    uint32_t box[2] = {32, 64};
    r = tcmap::encode(&m, p, 2, dims, strides, box, kind == 3);
  } else if (kind == 0 || kind == 4) {  // 32 px pitch: columns >= 26 are zero-filled
    uint64_t dims[4] = {32, 26, 26, (uint64_t)batch}, strides[3] = {128, 26 * 128, 676 * 128};
    uint32_t box[4] = {32, P_PITCH, 1, 1};  // one patch row per request
    r = tcmap::encode(&m, p, 4, dims, strides, box, kind == 4);
  } else {
    uint64_t dims[4] = {64, 24, 24, (uint64_t)batch}, strides[3] = {256, 24 * 256, 576 * 256};
    uint32_t box[4] = {32, P_PITCH, 1, 1};
    r = tcmap::encode(&m, p, 4, dims, strides, box, kind == 5);
  }
  if (r != B200_OK) return r;
  g_maps[key] = m;
  *out = m;
  return B200_OK;
}

}  // namespace

extern "C" int b200_conv2_fwd_tc2(const float* a1, const float* w, const float* bias, const void* ctrl, float* pooled,
                                  uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_fwd_tc2: batch must be > 0");
  CUtensorMap mp, mw;
  int r = get_map(a1, batch, 0, &mp);
  if (r != B200_OK) return r;
  r = get_map(w, 0, 1, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, V2_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  conv2_fwd_tc2_kernel<<<grid, 192, V2_SMEM, (cudaStream_t)stream>>>(mp, mw, bias, (const B200Ctrl*)ctrl, pooled, amax,
                                                                    nitems);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv2_dgrad_tc2(const float* dy, const float* w, const float* a1, float* da1, int batch,
                                    void* stream) {
  B200_REQUIRE(batch > 0, "conv2_dgrad_tc2: batch must be > 0");
  CUtensorMap md, mw;
  int r = get_map(dy, batch, 2, &md);
  if (r != B200_OK) return r;
  r = get_map(w, 0, 3, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  conv2_dgrad_tc2_kernel<<<grid, 192, DG_SMEM, (cudaStream_t)stream>>>(md, mw, a1, da1, nitems);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv2_wgrad_tc2(const float* dy, const float* a1, float* part, float* dw, int batch,
                                    void* stream) {
  B200_REQUIRE(batch > 0, "conv2_wgrad_tc2: batch must be > 0");
  cudaStream_t s = (cudaStream_t)stream;
  CUtensorMap ma, md;
  int r = get_map(a1, batch, 4, &ma);
  if (r != B200_OK) return r;
  r = get_map(dy, batch, 5, &md);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_wgrad_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    attr = true;
  }
  const int nitems = 3 * batch;
  const int grid = min(nitems, b200_sm_count());
  B200_CHECK(b200_launch(conv2_wgrad_tc2_kernel, dim3(grid), dim3(192), (size_t)WG_SMEM, s, ma, md, part, nitems));
  b200_count_launch();
  if (dw) b200_launch_reduce_partials(dw, part, 9 * C2 * C1, grid, CONV2_PART, s);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
