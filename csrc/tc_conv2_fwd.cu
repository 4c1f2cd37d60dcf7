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
// One kernel template, two producers (MMA issue loop and epilogue are shared):
//   FUSED = false : a1 comes from memory, patch rows loaded by TMA                (b200_conv2_fwd_tc)
//   FUSED = true  : conv1 is computed INSIDE the producer warps -- 9 FMAs per output from the uint8
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

B200_STAMP_ARRAY(g_stamps_fwd);
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
  // dropout bits of this thread's HB channels: the pooled pixel's 64 channels are one 128-bit Philox block
  uint32_t keep = 0xffffffffu;
  if (train & 2) {
    const uint4 r = dropout_block(seed, step, 1u, e_pixel >> 6);
    keep = block_word(r, (uint32_t)c0 >> 4) >> (2 * ((uint32_t)c0 & 15u));  // 2 bits per channel, from channel c0 on
  }
#pragma unroll
  for (int g4 = 0; g4 < HB / 4; ++g4) {
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      o[c] = fmaxf(m2[g4 * 4 + c] + bias_s[c0 + g4 * 4 + c], 0.f);
      if (train & 2) o[c] = ((keep >> (2 * (g4 * 4 + c))) & 3u) ? o[c] * (1.0f / (1.0f - P_DROP1)) : 0.f;
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

__device__ __forceinline__ float normalize_px4(uint8_t v) {
  // same operation order as ToTensor (div 255) then Normalize (sub mean, div std) -- see conv1.cu
  return __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), MEAN), STD);
}

// One kernel, two producers.
//   FUSED = false: a1 comes from memory; one producer warp loads the patch rows with TMA     (b200_conv2_fwd_tc)
//   FUSED = true : conv1 inside the producer.  Four producer warps build the 26-row x 16-pixel x 32-channel a1
//                  patch of an item directly in shared memory -- the same swizzled image TMA would leave,
//                      offset = r*2048 + px*128 + ((cq ^ (px & 7)) << 4)        (cq = channel quad)
//                  -- from the uint8 image (index gather, ToTensor + Normalize through a 256-entry table, 3x3 conv,
//                  bias, ReLU), publish it to the tensor core with fence.proxy.async + an mbarrier arrive per
//                  warp, and store it to global memory once for the backward pass.            (b200_conv12_fwd_tc)
// EW = number of epilogue warps: 8 (32 channels per warp) or 16 (16 channels per warp).
// Warp ids: 0..EW-1 epilogue (id % 4 = TMEM sub-partition), then the producer warp(s), last the MMA issuer (the warp
// arbiter favours high ids).
template <int EW, bool FUSED>
__global__ void __launch_bounds__((EW + (FUSED ? 4 : 1) + 1) * 32, 1)
conv_fwd_tc_kernel(const __grid_constant__ CUtensorMap mapP, const __grid_constant__ CUtensorMap mapW,
                   const uint8_t* __restrict__ images, const int32_t* __restrict__ idx,
                   const float* __restrict__ w1, const float* __restrict__ b1v,   // conv1 [9][32], [32]   (FUSED)
                   float* __restrict__ a1out,                                      // [B][26][26][32]       (FUSED)
                   const float* __restrict__ bias, const B200Ctrl* __restrict__ ctrl, float* __restrict__ pooled,
                   uint8_t* __restrict__ amax, int nitems) {
  constexpr int NP = FUSED ? 4 : 1;  // producer warps
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* wsm = smem;                        // [kh][kw][64 o][32 i]: per kh a K-major 192 x 32 B operand
  uint8_t* patch = smem + 9 * TAP_BYTES;      // [2][26 rows][16 px][128 B]
  __shared__ __align__(8) uint64_t w_bar, pfull[2], pempty[2], tfull[2], tempty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float bias_s[C2];
  __shared__ float xs[FUSED ? 2 : 1][FUSED ? IMG * IMG : 1];  // the item's normalised image, double buffered
  __shared__ float lut[FUSED ? 256 : 1];                      // uint8 -> ToTensor + Normalize (conv1_fwd_kernel's rounding)

  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_epi = warp_id < EW, is_mma = warp_id == EW + NP;
  B200_STAMP(g_stamps_fwd, 0);  // slot 0: kernel entry
  pdl_trigger();
  if (threadIdx.x == 0) {
    if (!FUSED) tma_prefetch_desc(&mapP);
    tma_prefetch_desc(&mapW);
    mbar_init(&w_bar, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&pfull[i], FUSED ? NP : 1); mbar_init(&pempty[i], 1);
      mbar_init(&tfull[i], 1); mbar_init(&tempty[i], EW);
    }
    fence_barrier_init();
  }
  if (is_mma) tmem_alloc(&tmem_base_smem, 512);  // 2 accumulator slots of 256 columns (192 used)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();  // on-chip set-up above overlaps the previous kernel's tail
  if (threadIdx.x < C2) bias_s[threadIdx.x] = bias[threadIdx.x];
  if (FUSED && threadIdx.x < 256) lut[threadIdx.x] = normalize_px4((uint8_t)threadIdx.x);
  __syncthreads();
  if (threadIdx.x == 0) B200_STAMP(g_stamps_fwd, 1);  // set-up done
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (!is_epi && !is_mma) {
    const int ptid = threadIdx.x - EW * 32;  // 0 .. 32*NP-1
    if (ptid < 32) {  // conv2 weights: once, by TMA
      const bool leader = elect_one();
      if (leader) mbar_expect_tx(&w_bar, 9 * TAP_BYTES);
      for (int t = 0; t < 9; ++t)
        if (leader) tma_load_2d(wsm + t * TAP_BYTES, &mapW, &w_bar, 0, t * 64);
    }
    if (!FUSED) {
      const bool leader = elect_one();
      for (int it = 0; it < my_items; ++it) {
        const int item = blockIdx.x + it * gridDim.x, n = item >> 1, X0 = (item & 1) * 10, buf = it & 1;
        mbar_wait(&pempty[buf], ((it >> 1) & 1) ^ 1);
        if (leader) mbar_expect_tx(&pfull[buf], F_PATCH_BYTES);
        for (int r = 0; r < F_ROWS; ++r)
          if (leader) tma_load_4d(patch + buf * F_PATCH_BYTES + r * F_ROW_BYTES, &mapP, &pfull[buf], 0, X0, r, n);
      }
    } else {
      // A thread owns one channel quad of one patch column and walks down the rows two at a time: 12 image values
      // from shared memory feed 72 FMAs, and its 36 conv1 weights stay in registers.
      const int cq = ptid & 7, px = (ptid >> 3) & 15;
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
      // The uint8 image of the NEXT item (index look-up, then 784 bytes) is fetched while the current one is
      // computed; the sample index is looked up one item further ahead still, so no load waits for an address.
      constexpr int NQ = IMG * IMG / 4;  // 196 uchar4 per image: threads 0..127 take one, threads 0..67 a second one
      const int pos0 = ctrl->pos;
      auto sample_of_item = [&](int it) -> int {
        if (it >= my_items) return 0;
        const int n = (blockIdx.x + it * gridDim.x) >> 1;
        return idx ? idx[pos0 + n] : pos0 + n;
      };
      uchar4 pf0 = make_uchar4(0, 0, 0, 0), pf1 = make_uchar4(0, 0, 0, 0);
      auto fetch = [&](int it, int sample) {
        if (it >= my_items) return;
        const uchar4* src = reinterpret_cast<const uchar4*>(images + (size_t)sample * (IMG * IMG));
        pf0 = src[ptid];
        if (ptid + 128 < NQ) pf1 = src[ptid + 128];
      };
      fetch(0, sample_of_item(0));
      int sample_next = sample_of_item(1);
      for (int it = 0; it < my_items; ++it) {
        const int item = blockIdx.x + it * gridDim.x, n = item >> 1, half = item & 1, X0 = half * 10, buf = it & 1;
        float* xb = xs[it & 1];  // double buffered: the previous item's readers are fenced off by its own barrier
        xb[ptid * 4 + 0] = lut[pf0.x]; xb[ptid * 4 + 1] = lut[pf0.y]; xb[ptid * 4 + 2] = lut[pf0.z]; xb[ptid * 4 + 3] = lut[pf0.w];
        if (ptid + 128 < NQ) {
          const int q = ptid + 128;
          xb[q * 4 + 0] = lut[pf1.x]; xb[q * 4 + 1] = lut[pf1.y]; xb[q * 4 + 2] = lut[pf1.z]; xb[q * 4 + 3] = lut[pf1.w];
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (ptid == 0) B200_STAMP(g_stamps_fwd, 2 + 3 * it);  // image of item `it` in shared memory
        fetch(it + 1, sample_next);
        sample_next = sample_of_item(it + 2);
        mbar_wait(&pempty[buf], ((it >> 1) & 1) ^ 1);
        if (ptid == 0) B200_STAMP(g_stamps_fwd, 3 + 3 * it);  // patch slot free
        uint8_t* pb = patch + buf * F_PATCH_BYTES;
        float* outg = a1out + (size_t)n * (H1 * W1 * C1);
        const int col = X0 + px;
        // the halves overlap in columns 10..15: half 0 owns columns 0..12, half 1 columns 13..25
        const bool owned = half ? (px >= 3) : (px <= 12);
        uint8_t* pcol = pb + px * 128 + ((cq ^ (px & 7)) << 4);
#pragma unroll 1
        for (int r = 0; r < F_ROWS; r += 2) {
          float x[4][3];
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) x[a][b] = xb[(r + a) * IMG + col + b];
          float acc[2][4];
#pragma unroll
          for (int c = 0; c < 4; ++c) { acc[0][c] = br[c]; acc[1][c] = br[c]; }
          // same FMA order per output as conv1_fwd_kernel (conv1.cu): bias, then taps kh-major
#pragma unroll
          for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw)
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                acc[0][c] = fmaf(x[kh][kw], wr[kh * 3 + kw][c], acc[0][c]);
                acc[1][c] = fmaf(x[kh + 1][kw], wr[kh * 3 + kw][c], acc[1][c]);
              }
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            const float4 o = make_float4(fmaxf(acc[rr][0], 0.f), fmaxf(acc[rr][1], 0.f), fmaxf(acc[rr][2], 0.f),
                                         fmaxf(acc[rr][3], 0.f));
            *reinterpret_cast<float4*>(pcol + (r + rr) * F_ROW_BYTES) = o;
            if (owned) *reinterpret_cast<float4*>(outg + ((size_t)(r + rr) * W1 + col) * C1 + cq * 4) = o;
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&pfull[buf]);
        if (ptid == 0) B200_STAMP(g_stamps_fwd, 4 + 3 * it);  // patch of item `it` published
      }
    }
  } else if (is_mma) {
    const bool leader = elect_one();
    constexpr uint32_t idesc = idesc_tf32(128, 192, 0, 0);
    mbar_wait(&w_bar, 0);
    B200_STAMP(g_stamps_fwd, 14);  // conv2 weights in shared memory
    int tc_ = 0;
    for (int it = 0; it < my_items; ++it) {
      const int buf = it & 1;
      mbar_wait(&pfull[buf], (it >> 1) & 1);
      tc_fence_after();
      B200_STAMP(g_stamps_fwd, 15 + it);  // MMA warp sees the patch of item `it`
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
    const int q = warp_id & 3, oc = warp_id >> 2;
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
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + slot * 256 + oc * NCH;
        const int ph = ty * 4 + q, pw = (X0 + lx) >> 1;
        const uint32_t e_pixel = ((uint32_t)(n * HP + ph) * WP + pw) * C2;
        float v0[NCH], v1[NCH], v2[NCH];
        tmem_ld_nowait<NCH>(taddr, v0);        // kw = 0
        tmem_ld_nowait<NCH>(taddr + 64, v1);   // kw = 1
        tmem_ld_nowait<NCH>(taddr + 128, v2);  // kw = 2
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[slot]);
        if (warp_id == 0) B200_STAMP(g_stamps_fwd, 20 + 2 * tc_);  // accumulator of tile tc_ in registers
#pragma unroll
        for (int j = 0; j < NCH; ++j)
          v0[j] += __shfl_down_sync(0xffffffffu, v1[j], 1) + __shfl_down_sync(0xffffffffu, v2[j], 2);
        pool_store<NCH>(v0, lane, oc, bias_s, train, seed, step, e_pixel, store, pooled, amax);
        if (warp_id == 0) B200_STAMP(g_stamps_fwd, 21 + 2 * tc_);  // tile tc_ stored
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (is_mma) tmem_dealloc(tmem_base, 512);
  if (threadIdx.x == 0) B200_STAMP(g_stamps_fwd, 47);  // kernel exit
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

template <int EW, bool FUSED>
int launch_fwd(const CUtensorMap& mp, const CUtensorMap& mw, const uint8_t* images, const int32_t* idx, const float* w1,
               const float* b1, float* a1, const float* bias, const void* ctrl, float* pooled, uint8_t* amax, int batch,
               cudaStream_t s) {
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv_fwd_tc_kernel<EW, FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  const dim3 block((EW + (FUSED ? 4 : 1) + 1) * 32);
  // the fused kernel is the first of a training step: it follows copies, other streams' joins and the previous
  // step's end marker, and is launched with full stream serialisation
  if (FUSED)
    B200_CHECK(b200_launch_serial(conv_fwd_tc_kernel<EW, FUSED>, dim3(grid), block, (size_t)F3_SMEM, s, mp, mw, images, idx, w1, b1,
                                  a1, bias, (const B200Ctrl*)ctrl, pooled, amax, nitems));
  else
    B200_CHECK(b200_launch(conv_fwd_tc_kernel<EW, FUSED>, dim3(grid), block, (size_t)F3_SMEM, s, mp, mw, images, idx, w1, b1, a1,
                           bias, (const B200Ctrl*)ctrl, pooled, amax, nitems));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Epilogue warps (B200_FWD_EW = 8 | 16).  Measured, batch 200, us per training step: the plain conv2 kernel is
// ~1.5 us faster with 16 (more warps hide the shuffle / TMEM latencies); in the fused kernel 16 epilogue warps
// starve the four conv1 producer warps (patch of an item ready after 5.0 us instead of 3.5) and the step is 4.7 us
// SLOWER (124.0 vs 119.3), so the fused kernel defaults to 8.
int epilogue_warps(bool fused) {
  static const int ew = getenv("B200_FWD_EW") ? atoi(getenv("B200_FWD_EW")) : 0;
  return ew == 8 || ew == 16 ? ew : (fused ? 8 : 16);
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
  cudaStream_t s = (cudaStream_t)stream;
  return epilogue_warps(false) == 8
             ? launch_fwd<8, false>(mp, mw, nullptr, nullptr, nullptr, nullptr, nullptr, bias, ctrl, pooled, amax, batch, s)
             : launch_fwd<16, false>(mp, mw, nullptr, nullptr, nullptr, nullptr, nullptr, bias, ctrl, pooled, amax, batch, s);
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
  cudaStream_t s = (cudaStream_t)stream;
  return epilogue_warps(true) == 8
             ? launch_fwd<8, true>(mw, mw, images, idx, w1, b1, a1, b2, ctrl, pooled, amax, batch, s)
             : launch_fwd<16, true>(mw, mw, images, idx, w1, b1, a1, b2, ctrl, pooled, amax, batch, s);
}

#ifdef B200_STAMPS
extern "C" int b200_debug_stamps_fwd(unsigned long long* out_host) {
  B200_CHECK(cudaMemcpyFromSymbol(out_host, g_stamps_fwd, sizeof(unsigned long long) * 160 * 48));
  return B200_OK;
}
#endif
