// conv2 kernels, third generation: "kw-scatter" formulation to get off the shared-memory bound.
//
// Measured on B200 (profiles/README.md): an SS-mode 128 x N x 8 tf32 tcgen05.mma costs about
// (A bytes + B bytes) / 64 cycles — the operand read from shared memory, not the tensor pipe, sets
// the pace when N is small.  The v2 dgrad issues 72 MMAs of N = 32 per 128 output pixels (80 cycles
// each for 16 cycles of tensor work).  Here the three kw taps are moved from the A side to the N side:
//
//   D[pixel (y,x')][(kw,i)] = sum_kh sum_o dY[y-kh][x'][o] * W[kh][kw][o][i]          (N = 96)
//   da1[y][x][i]            = relu'(a1) * sum_kw D[(y, x-kw)][(kw,i)]
//
// so one 128 x 96 x 8 MMA replaces three 128 x 32 x 8 ones and re-uses the same A tile; the kh taps
// are still accumulated by the tensor core through row-shifted views of the dY patch; the kw sum
// becomes two warp shuffles per value in the epilogue, because a tile is 4 output rows x 32 pixels
// and a warp holds exactly one row (lane = x).  With the patch pitch padded to 32 pixels such a tile
// is a plain contiguous 16 KiB K-major operand.  24 MMAs per tile instead of 72.
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
constexpr int DG_SMEM = 9 * TAP_BYTES + 2 * PATCH_BYTES + 1024;

// Item = (image, half): half 0 -> output rows 0..15 (4 tiles), half 1 -> rows 16..25 (3 tiles, the last
// one half empty).  The patch holds dY rows y0-2 .. y0+15, columns 0..31 (24 real, zero beyond).
__global__ void __launch_bounds__(320, 1)
conv2_dgrad_tc3_kernel(const __grid_constant__ CUtensorMap mapDY, const __grid_constant__ CUtensorMap mapW,
                       const float* __restrict__ a1, float* __restrict__ da1, int nitems, int flags) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                     // [kh][kw][64 o][32 i], 32B-atom swizzle: N-major B operand
  uint8_t* patch = smem + 9 * TAP_BYTES;   // [2 channel chunks][18][32][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[4], tempty[4];
  __shared__ uint32_t tmem_base_smem;

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapDY);
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&pfull[i], 1); mbar_init(&pempty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 8); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, 512);  // 4 accumulator slots of 128 columns (96 used)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();  // on-chip set-up above overlaps the previous kernel's tail
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  // Items alternate between 4-tile (upper half) and 3-tile (lower half) work.  With an even grid a CTA
  // would see the same half in every round (12 tiles against 9); swapping neighbours in odd rounds gives
  // every CTA both kinds (11 against 10).  nitems and the grid are even, so pairs stay inside a round.
  const bool flip = (gridDim.x & 1) == 0 && (flags & 1);
  auto item_of = [&](int it) { return (int)(blockIdx.x + it * gridDim.x) ^ ((flip && (it & 1)) ? 1 : 0); };

  if (warp == 0) {
    const bool leader = elect_one();
    if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
    for (int t = 0; t < 9; ++t)
      if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
    for (int it = 0; it < my_items; ++it) {
      const int item = item_of(it), n = item >> 1, y0 = (item & 1) * 16;
      for (int kc = 0; kc < 2; ++kc) {
        mbar_wait(&pempty[kc], (it & 1) ^ 1);
        if (leader) mbar_expect_tx(&pfull[kc], PATCH_BYTES);
        for (int r = 0; r < P_ROWS; ++r)
          if (leader) tma_load_4d(patch + kc * PATCH_BYTES + r * ROW_BYTES, &mapDY, &pfull[kc], kc * 32, 0, y0 - 2 + r, n);
      }
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    constexpr uint32_t idesc = idesc_tf32(128, 96, 0, 1);
    mbar_wait(&w_bar, 0);
    int tc_ = 0;  // running tile counter -> TMEM slot
    for (int it = 0; it < my_items; ++it) {
      const int ntile = (item_of(it) & 1) ? 3 : 4;
      // chunk-major order: all tiles of the item consume channel chunk 0, then chunk 1
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
    // 8 epilogue warps: warp pairs (w, w+4) share a TMEM sub-partition and split the 32 channels
    const int q = wq, hc = (warp - 2) >> 2;
    int tc_ = 0;
    for (int it = 0; it < my_items; ++it) {
      const int item = item_of(it), n = item >> 1, half = item & 1, y0 = half * 16;
      const int ntile = half ? 3 : 4;
      for (int t = 0; t < ntile; ++t) {
        const int slot = (tc_ + t) & 3;
        // the ReLU mask (conv1's output) does not depend on the MMAs: fetch it before waiting for them
        const int y = y0 + 4 * t + q, x = lane;
        const bool valid = y < H1 && x < W1;
        const size_t e = (((size_t)n * H1 + (valid ? y : 0)) * W1 + (valid ? x : 0)) * C1 + hc * 16;
        float4 am[4];
        if (flags & 2) {
#pragma unroll
          for (int j = 0; j < 4; ++j) am[j] = *reinterpret_cast<const float4*>(a1 + e + 4 * j);
        }
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
        // da1[y][x] = D[x][kw=0] + D[x-1][kw=1] + D[x-2][kw=2]   (lane = x; columns >= 24 of D are zero)
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float s1 = __shfl_up_sync(0xffffffffu, v1[j], 1);
          const float s2 = __shfl_up_sync(0xffffffffu, v2[j], 2);
          v0[j] += (lane >= 1 ? s1 : 0.f) + (lane >= 2 ? s2 : 0.f);
        }
        if (!(flags & 2)) {
#pragma unroll
          for (int j = 0; j < 4; ++j) am[j] = *reinterpret_cast<const float4*>(a1 + e + 4 * j);
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
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// =============================================================================================
// forward v3:  D[(y,x')][(kw,o)] = sum_kh sum_i a1[y+kh][x'][i] * W[kh][kw][o][i]      (N = 192)
//              out[y][x][o]      = sum_kw D[(y, x+kw)][(kw,o)]   -> bias, ReLU, 2x2 max-pool, dropout
// Item = (image, x-half): a 26-row x 16-pixel patch (columns 0..15 or 10..25, 52 KB, pitch 16 pixels)
// feeds three 8-row x 16-pixel tiles; a tile is a contiguous 16 KiB K-major operand at row (8*ty + kh).
// The B operand of one kh is simply the native weight block [kw][o][i] = 192 rows of 128 bytes.
// 12 MMAs of 128x192x8 per tile instead of 36 of 128x64x8.  A warp holds two tile rows (lanes 0..15 =
// row 2q, lanes 16..31 = row 2q+1), so the kw-sum is a shuffle-down by 1 and 2 and the 2x2 pool is the
// (lane^1, lane^16) neighbourhood.  14 of the 16 columns are valid outputs; the two halves overlap.
// =============================================================================================
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
conv2_fwd_tc3_kernel(const __grid_constant__ CUtensorMap mapP, const __grid_constant__ CUtensorMap mapW,
                     const float* __restrict__ bias, const B200Ctrl* __restrict__ ctrl, float* __restrict__ pooled,
                     uint8_t* __restrict__ amax, int nitems, int dbg) {
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
            if (leader && !(dbg & 2)) umma_tf32(d, desc_kmajor(sa + kk * 32), desc_kmajor(sw + kk * 32), idesc, (kh | kk) != 0);
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
        if (dbg & 1) continue;
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

struct Key {
  const void* p; int batch, kind;
  bool operator<(const Key& o) const { return std::tie(p, batch, kind) < std::tie(o.p, o.batch, o.kind); }
};
std::map<Key, CUtensorMap> g_maps;
std::mutex g_mu;

// kind 0: dY patch row (K-major, 32 channels x 32 px x 1 row)   1: weights (32B-atom swizzle, N-major)
int get_map(const float* p, int batch, int kind, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  Key key{p, batch, kind};
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return B200_OK; }
  CUtensorMap m;
  int r;
  if (kind == 0) {
    uint64_t dims[4] = {64, 24, 24, (uint64_t)batch}, strides[3] = {256, 24 * 256, 576 * 256};
    uint32_t box[4] = {32, P_PITCH, 1, 1};
    r = tcmap::encode(&m, p, 4, dims, strides, box, 0);
  } else if (kind == 2) {  // a1 patch row for the forward: 32 channels x 16 px x 1 row
    uint64_t dims[4] = {32, 26, 26, (uint64_t)batch}, strides[3] = {128, 26 * 128, 676 * 128};
    uint32_t box[4] = {32, F_PITCH, 1, 1};
    r = tcmap::encode(&m, p, 4, dims, strides, box, 0);
  } else if (kind == 3) {  // weights, K-major (plain 128B swizzle)
    uint64_t dims[2] = {32, 576}, strides[1] = {128};
    uint32_t box[2] = {32, 64};
    r = tcmap::encode(&m, p, 2, dims, strides, box, 0);
  } else {
    uint64_t dims[2] = {32, 576}, strides[1] = {128};
    uint32_t box[2] = {32, 64};
    r = tcmap::encode(&m, p, 2, dims, strides, box, 1);
  }
  if (r != B200_OK) return r;
  g_maps[key] = m;
  *out = m;
  return B200_OK;
}

}  // namespace

extern "C" int b200_conv2_dgrad_tc3(const float* dy, const float* w, const float* a1, float* da1, int batch,
                                    void* stream) {
  B200_REQUIRE(batch > 0, "conv2_dgrad_tc3: batch must be > 0");
  CUtensorMap md, mw;
  int r = get_map(dy, batch, 0, &md);
  if (r != B200_OK) return r;
  r = get_map(w, 0, 1, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  static const int flags = getenv("B200_DG_FLAGS") ? atoi(getenv("B200_DG_FLAGS")) : 3;  // 1 balance items, 2 prefetch mask
  B200_CHECK(b200_launch(conv2_dgrad_tc3_kernel, dim3(grid), dim3(320), (size_t)DG_SMEM, (cudaStream_t)stream, md, mw, a1,
                         da1, nitems, flags));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_conv2_fwd_tc3(const float* a1, const float* w, const float* bias, const void* ctrl, float* pooled,
                                  uint8_t* amax, int batch, void* stream) {
  B200_REQUIRE(batch > 0, "conv2_fwd_tc3: batch must be > 0");
  CUtensorMap mp, mw;
  int r = get_map(a1, batch, 2, &mp);
  if (r != B200_OK) return r;
  r = get_map(w, 0, 3, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_tc3_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    B200_CHECK(cudaFuncSetAttribute(conv2_fwd_tc3_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  static const int ew = getenv("B200_FWD_EW") ? atoi(getenv("B200_FWD_EW")) : 16;
  static const int dbg = getenv("B200_DBG") ? atoi(getenv("B200_DBG")) : 0;
  if (ew == 8)
    B200_CHECK(b200_launch(conv2_fwd_tc3_kernel<8>, dim3(grid), dim3(320), (size_t)F3_SMEM, (cudaStream_t)stream, mp, mw, bias,
                           (const B200Ctrl*)ctrl, pooled, amax, nitems, dbg));
  else
    B200_CHECK(b200_launch(conv2_fwd_tc3_kernel<16>, dim3(grid), dim3(576), (size_t)F3_SMEM, (cudaStream_t)stream, mp, mw, bias,
                           (const B200Ctrl*)ctrl, pooled, amax, nitems, dbg));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
