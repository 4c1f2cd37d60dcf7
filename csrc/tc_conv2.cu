// conv2 (32 -> 64 channels, 3x3, no padding) as implicit GEMMs on tcgen05 tensor cores.
// Replaces cuDNN's conv forward / backward-data / backward-filter plus the separate bias, ReLU,
// max-pool and dropout kernels (mnist_ddp.py:52-55 and its backward; SURVEY N2-N5, N12, N13,
// section 8 rows 3, 8, 9).
//
// Activations are NHWC with C = 32 fp32 = 128 bytes, i.e. ONE PIXEL IS ONE 128-BYTE SWIZZLE ROW:
// a TMA box over (c, w, h, n) lands in shared memory directly as a tcgen05 operand tile, and a
// 3x3 tap is just a shifted box — no im2col buffer, no smem gather.  The same box is a K-major
// [pixels][32 channels] tile (forward, dgrad: K = channels) or an MN-major [pixels][32 channels]
// tile (wgrad: K = pixels).
//
//   forward : persistent CTAs, tile = 8x8 output pixels of 2 images (128 GEMM rows) x 64 channels;
//             9 taps x 4 tf32 MMAs (K = 8) accumulate in TMEM (double buffered); the epilogue warps
//             do 2x2 max-pool with warp shuffles straight out of tcgen05.ld registers, then bias,
//             ReLU, Philox dropout, and store the pooled NHWC tensor + 2-bit argmax.
//   dgrad   : tile = 26 px x 2 rows x 2 images (104 of 128 rows) x 32 channels from the dense dY;
//             negative / overflowing tap coordinates are zero-filled by TMA; epilogue applies conv1's
//             ReLU mask.
//   wgrad   : K = pixels.  D[(tap,i)][o] += a1(tap-shifted)[pix][i]^T . dY[pix][o]; three M = 128
//             accumulators (4 taps each) stay in TMEM over the CTA's whole pixel range; each CTA
//             writes one partial, a fixed-order reduction follows.
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "tc_common.cuh"

using namespace tc;
using namespace net;

namespace {

constexpr int W2_ELEMS = 9 * C2 * C1;
constexpr int TAP_BYTES = 64 * 128;  // one tap of weights, or one 8x8-pixel box of 32 channels: 8 KiB

// =============================================================================================
// forward
// =============================================================================================
constexpr int F_STAGES = 6;
constexpr int F_A_BYTES = 128 * 128;  // 2 images x 64 pixels x 128 B
constexpr int F_SMEM = 9 * TAP_BYTES + F_STAGES * F_A_BYTES + 1024;

__global__ void __launch_bounds__(192, 1)
conv2_fwd_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapW,
                    const float* __restrict__ bias, const B200Ctrl* __restrict__ ctrl, float* __restrict__ pooled,
                    uint8_t* __restrict__ amax, int batch, int ntiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                       // [9][64 o][32 i]  K-major B operand per tap
  uint8_t* asm_ = smem + 9 * TAP_BYTES;      // [F_STAGES][128 rows][128 B]
  __shared__ __align__(8) uint64_t full_bar[F_STAGES], empty_bar[F_STAGES], w_bar, tfull_bar[2], tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[C2];

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA);
    tma_prefetch_desc(&mapW);
    for (int s = 0; s < F_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&w_bar, 1);
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 4); }
    fence_barrier_init();
  }
  if (threadIdx.x < C2) bias_s[threadIdx.x] = bias[threadIdx.x];
  if (warp == 1) tmem_alloc(&tmem_base_smem, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t) if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int tile = blockIdx.x + it * gridDim.x;
        const int pair = tile / 9, blk = tile - pair * 9, by = (blk / 3) * 8, bx = (blk % 3) * 8;
        for (int tap = 0; tap < 9; ++tap, ++g) {
          const int s = g % F_STAGES;
          mbar_wait(&empty_bar[s], ((g / F_STAGES) & 1) ^ 1);
          if (leader) mbar_expect_tx(&full_bar[s], F_A_BYTES);
          if (leader) tma_load_4d(asm_ + s * F_A_BYTES, &mapA, &full_bar[s], 0, bx + tap % 3, by + tap / 3, pair * 2);
        }
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(128, 64, 0, 0);
      mbar_wait(&w_bar, 0);
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int acc = it & 1;
        mbar_wait(&tempty_bar[acc], ((it >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d = tmem_base + acc * 64;
        for (int tap = 0; tap < 9; ++tap, ++g) {
          const int s = g % F_STAGES;
          mbar_wait(&full_bar[s], (g / F_STAGES) & 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(asm_ + s * F_A_BYTES), sw = smem_u32(wsm + tap * TAP_BYTES);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            if (leader) umma_tf32(d, desc_kmajor(sa + kk * 32), desc_kmajor(sw + kk * 32), idesc, (tap | kk) != 0);
          if (leader) umma_commit(&empty_bar[s]);
        }
        if (leader) umma_commit(&tfull_bar[acc]);
      }
    }
  } else {
    const int q = wq;
    const int train = ctrl->train;
    const unsigned long long seed = ctrl->seed, step = ctrl->step;
    const bool b0 = lane & 1, b3 = (lane >> 3) & 1;
    const int py = (q & 1) * 4 + (lane >> 3), px = lane & 7;
    const int c0 = (b0 ? 32 : 0) + (b3 ? 16 : 0);
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = blockIdx.x + it * gridDim.x;
      const int pair = tile / 9, blk = tile - pair * 9, by = (blk / 3) * 8, bx = (blk % 3) * 8;
      const int acc = it & 1;
      mbar_wait(&tfull_bar[acc], (it >> 1) & 1);
      tc_fence_after();
      float v0[32], v1[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * 64;
      tmem_ld32(taddr, v0);
      tmem_ld32(taddr + 32, v1);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);  // TMEM buffer may be overwritten from here on

      // ---- 2x2 max-pool across lanes: x-neighbour = lane^1, y-neighbour = lane^8.  Each exchange
      // halves the channels a lane keeps, so the 4 lanes of a window end with 16 channels each.
      float m1[32];
      uint32_t dxmask = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float mine = b0 ? v1[j] : v0[j], send = b0 ? v0[j] : v1[j];
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        const float ev = b0 ? recv : mine, od = b0 ? mine : recv;  // even / odd column of the window
        const bool gt = od > ev;                                     // strict: first maximum wins
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
        const float top = b3 ? recv : mine, bot = b3 ? mine : recv;  // even / odd row of the window
        const uint32_t dxt = b3 ? dxo : dxm, dxb = b3 ? dxm : dxo;
        const bool gt = bot > top;
        m2[j] = gt ? bot : top;
        am[j] = gt ? (2u + dxb) : dxt;
      }
      const int n = pair * 2 + (q >> 1);
      if (n < batch) {
        const int ph = (by + py) >> 1, pw = (bx + px) >> 1;
        const uint32_t e = ((uint32_t)(n * HP + ph) * WP + pw) * C2 + c0;
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
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 128);
}

// =============================================================================================
// dgrad
// =============================================================================================
constexpr int D_STAGES = 6;
constexpr int D_A_BYTES = 104 * 128;   // 26 x 2 x 2 pixels
constexpr int D_A_STRIDE = 128 * 128;  // stage pitch (the MMA reads 128 rows; rows 104.. are don't-care)
constexpr int D_SMEM = 9 * TAP_BYTES + D_STAGES * D_A_STRIDE + 1024;

__global__ void __launch_bounds__(192, 1)
conv2_dgrad_tc_kernel(const __grid_constant__ CUtensorMap mapDY, const __grid_constant__ CUtensorMap mapW,
                      const float* __restrict__ a1, float* __restrict__ da1, int batch, int ntiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                    // [9][64 o][32 i]: N-major B operand (K = o rows)
  uint8_t* asm_ = smem + 9 * TAP_BYTES;
  __shared__ __align__(8) uint64_t full_bar[D_STAGES], empty_bar[D_STAGES], w_bar, tfull_bar[2], tempty_bar[2];
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
    for (int s = 0; s < D_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&w_bar, 1);
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 4); }
    fence_barrier_init();
  }
  // rows 104..127 of every stage are never written by TMA: clear them once so the MMA reads zeros
  for (int e = threadIdx.x; e < D_STAGES * (24 * 128 / 16); e += blockDim.x) {
    const int s = e / (24 * 128 / 16), r = e % (24 * 128 / 16);
    reinterpret_cast<uint4*>(asm_ + s * D_A_STRIDE + D_A_BYTES)[r] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  if (warp == 1) tmem_alloc(&tmem_base_smem, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t) if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int tile = blockIdx.x + it * gridDim.x;
        const int pair = tile / 13, y0 = (tile - pair * 13) * 2;
        for (int kb = 0; kb < 18; ++kb, ++g) {
          const int tap = kb >> 1, kc = kb & 1, kh = tap / 3, kw = tap - kh * 3;
          const int s = g % D_STAGES;
          mbar_wait(&empty_bar[s], ((g / D_STAGES) & 1) ^ 1);
          if (leader) mbar_expect_tx(&full_bar[s], D_A_BYTES);
          if (leader) tma_load_4d(asm_ + s * D_A_STRIDE, &mapDY, &full_bar[s], kc * 32, -kw, y0 - kh, pair * 2);
        }
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(128, 32, 0, 1);
      mbar_wait(&w_bar, 0);
      int g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int acc = it & 1;
        mbar_wait(&tempty_bar[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc * 32;
        for (int kb = 0; kb < 18; ++kb, ++g) {
          const int tap = kb >> 1, kc = kb & 1;
          const int s = g % D_STAGES;
          mbar_wait(&full_bar[s], (g / D_STAGES) & 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(asm_ + s * D_A_STRIDE);
          const uint32_t sw = smem_u32(wsm + tap * TAP_BYTES + kc * 4096);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            if (leader) umma_tf32(d, desc_kmajor(sa + kk * 32), desc_mnmajor(sw + kk * 1024, TAP_BYTES), idesc, (kb | kk) != 0);
          if (leader) umma_commit(&empty_bar[s]);
        }
        if (leader) umma_commit(&tfull_bar[acc]);
      }
    }
  } else {
    const int q = wq;
    const int r = q * 32 + lane;
    const int nl = r / 52, rem = r - nl * 52, hy = rem / 26, x = rem - hy * 26;
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = blockIdx.x + it * gridDim.x;
      const int pair = tile / 13, y0 = (tile - pair * 13) * 2;
      const int acc = it & 1;
      mbar_wait(&tfull_bar[acc], (it >> 1) & 1);
      tc_fence_after();
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * 32, v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      const int n = pair * 2 + nl;
      if (r < 104 && n < batch) {
        const size_t e = (((size_t)n * H1 + y0 + hy) * W1 + x) * C1;
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
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 64);
}

// =============================================================================================
// wgrad
// =============================================================================================
constexpr int G_STAGES = 2;
constexpr int G_STAGE_BYTES = 12 * TAP_BYTES;  // 9 tap-shifted a1 boxes, 2 dY channel chunks, 1 pad slot
constexpr int G_LOAD_BYTES = 11 * TAP_BYTES;
constexpr int G_SMEM = G_STAGES * G_STAGE_BYTES + 1024;
constexpr int CONV2_PART = W2_ELEMS + C2;

__global__ void __launch_bounds__(192, 1)
conv2_wgrad_tc_kernel(const __grid_constant__ CUtensorMap mapA1, const __grid_constant__ CUtensorMap mapDY,
                      float* __restrict__ part, int nkblk) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[G_STAGES], empty_bar[G_STAGES], tfull_bar;
  __shared__ uint32_t tmem_base_smem;

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA1);
    tma_prefetch_desc(&mapDY);
    for (int s = 0; s < G_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tfull_bar, 1);
    fence_barrier_init();
  }
  // the pad slot of each stage is read (as don't-care rows) by the third MMA: keep it finite
  for (int e = threadIdx.x; e < G_STAGES * (TAP_BYTES / 16); e += blockDim.x) {
    const int s = e / (TAP_BYTES / 16), r = e % (TAP_BYTES / 16);
    reinterpret_cast<uint4*>(smem + s * G_STAGE_BYTES + G_LOAD_BYTES)[r] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  if (warp == 1) tmem_alloc(&tmem_base_smem, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int my_blocks = (nkblk - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      for (int it = 0; it < my_blocks; ++it) {
        const int kblk = blockIdx.x + it * gridDim.x;
        const int n = kblk / 9, blk = kblk - n * 9, by = (blk / 3) * 8, bx = (blk % 3) * 8;
        const int s = it % G_STAGES;
        mbar_wait(&empty_bar[s], ((it / G_STAGES) & 1) ^ 1);
        uint8_t* st = smem + s * G_STAGE_BYTES;
        if (leader) mbar_expect_tx(&full_bar[s], G_LOAD_BYTES);
        for (int tap = 0; tap < 9; ++tap)
          if (leader) tma_load_4d(st + tap * TAP_BYTES, &mapA1, &full_bar[s], 0, bx + tap % 3, by + tap / 3, n);
        if (leader) tma_load_4d(st + 9 * TAP_BYTES, &mapDY, &full_bar[s], 0, bx, by, n);
        if (leader) tma_load_4d(st + 10 * TAP_BYTES, &mapDY, &full_bar[s], 32, bx, by, n);
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(128, 64, 1, 1);
      for (int it = 0; it < my_blocks; ++it) {
        const int s = it % G_STAGES;
        mbar_wait(&full_bar[s], (it / G_STAGES) & 1);
        tc_fence_after();
        const uint32_t st = smem_u32(smem + s * G_STAGE_BYTES);
#pragma unroll 1
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t bd = desc_mnmajor(st + 9 * TAP_BYTES + kk * 1024, TAP_BYTES);
#pragma unroll
          for (int grp = 0; grp < 3; ++grp)
            if (leader) umma_tf32(tmem_base + grp * 64, desc_mnmajor(st + grp * 4 * TAP_BYTES + kk * 1024, TAP_BYTES), bd, idesc,
                      (it | kk) != 0);
        }
        if (leader) umma_commit(&empty_bar[s]);
      }
      if (leader) umma_commit(&tfull_bar);
    }
  } else {
    const int q = wq;
    mbar_wait(&tfull_bar, 0);
    tc_fence_after();
    float* dst = part + (size_t)blockIdx.x * CONV2_PART;
    for (int grp = 0; grp < 3; ++grp) {
      const int tap = grp * 4 + q;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + grp * 64 + c0, v);
        if (tap < 9) {
          if (my_blocks == 0) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0.f;
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) dst[(tap * C2 + c0 + j) * C1 + lane] = v[j];  // [tap][o][i], i = lane
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// ---- host ---------------------------------------------------------------------------------------
struct Key {
  const void* p; int batch, kind;
  bool operator<(const Key& o) const { return std::tie(p, batch, kind) < std::tie(o.p, o.batch, o.kind); }
};
std::map<Key, CUtensorMap> g_maps;
std::mutex g_mu;

// kind: 0 a1 K-major box 8x8x2 | 1 weights K-major | 2 dY K-major box 26x2x2 | 3 weights MN-major
//       4 a1 MN-major box 8x8x1 | 5 dY MN-major box 8x8x1
int get_map(const float* p, int batch, int kind, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  Key key{p, batch, kind};
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return B200_OK; }
  CUtensorMap m;
  int r;
  if (kind == 1 || kind == 3) {
    uint64_t dims[2] = {32, 576}, strides[1] = {128};
    uint32_t box[2] = {32, 64};
    r = tcmap::encode(&m, p, 2, dims, strides, box, kind == 3);
  } else if (kind == 0 || kind == 4) {
    uint64_t dims[4] = {32, 26, 26, (uint64_t)batch}, strides[3] = {128, 26 * 128, 676 * 128};
    uint32_t box[4] = {32, 8, 8, (uint32_t)(kind == 0 ? 2 : 1)};
    r = tcmap::encode(&m, p, 4, dims, strides, box, kind == 4);
  } else {
    uint64_t dims[4] = {64, 24, 24, (uint64_t)batch}, strides[3] = {256, 24 * 256, 576 * 256};
    uint32_t box2[4] = {32, 26, 2, 2}, box5[4] = {32, 8, 8, 1};
    r = tcmap::encode(&m, p, 4, dims, strides, kind == 2 ? box2 : box5, kind == 5);
  }
  if (r != B200_OK) return r;
  g_maps[key] = m;
  *out = m;
  return B200_OK;
}

#define TRY(x)                    \
  do {                            \
    int _r = (x);                 \
    if (_r != B200_OK) return _r; \
  } while (0)

}  // namespace

extern "C" int b200_conv2_fwd_tc(const float* a1, const float* w, const float* bias, const void* ctrl, float* pooled,
                                 uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_fwd_tc: batch must be > 0");
  CUtensorMap ma, mw;
  TRY(get_map(a1, batch, 0, &ma));
  TRY(get_map(w, 0, 1, &mw));
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM));
    attr = true;
  }
  const int ntiles = 9 * ceil_div(batch, 2);
  const int grid = min(ntiles, b200_sm_count());
  conv2_fwd_tc_kernel<<<grid, 192, F_SMEM, (cudaStream_t)stream>>>(ma, mw, bias, (const B200Ctrl*)ctrl, pooled, amax,
                                                                  batch, ntiles);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv2_dgrad_tc(const float* dy, const float* w, const float* a1, float* da1, int batch,
                                   void* stream) {
  B200_REQUIRE(batch > 0, "conv2_dgrad_tc: batch must be > 0");
  CUtensorMap md, mw;
  TRY(get_map(dy, batch, 2, &md));
  TRY(get_map(w, 0, 3, &mw));
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D_SMEM));
    attr = true;
  }
  const int ntiles = 13 * ceil_div(batch, 2);
  const int grid = min(ntiles, b200_sm_count());
  conv2_dgrad_tc_kernel<<<grid, 192, D_SMEM, (cudaStream_t)stream>>>(md, mw, a1, da1, batch, ntiles);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// the bias gradient comes from b200_unpool_dense (column sums of the pooled gradient)
extern "C" int b200_conv2_wgrad_tc(const float* dy, const float* a1, float* part, float* dw, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_wgrad_tc: batch must be > 0");
  cudaStream_t s = (cudaStream_t)stream;
  CUtensorMap ma, md;
  TRY(get_map(a1, batch, 4, &ma));
  TRY(get_map(dy, batch, 5, &md));
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM));
    attr = true;
  }
  const int nkblk = 9 * batch;
  const int grid = min(nkblk, b200_sm_count());
  conv2_wgrad_tc_kernel<<<grid, 192, G_SMEM, s>>>(ma, md, part, nkblk);
  b200_count_launch();
  if (dw) b200_launch_reduce_partials(dw, part, W2_ELEMS, grid, CONV2_PART, s);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
