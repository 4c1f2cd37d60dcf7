// conv2 backward on the 5th-generation tensor cores, without a dense dY tensor
// (mnist_ddp.py:52-55 backward inside :72).
//
// Max-pool backward is "the value goes to the arg-max position of its 2x2 window, zeros elsewhere", so the
// 29.5 MB (604 MB at batch 4096) gradient dY[B][24][24][64] of conv2's output is fully described by the pooled
// gradient dp[B][12][12][64] (already masked by dropout/ReLU in the fc1 dgrad epilogue) and the 2-bit arg-max
// codes the forward stored.  Both kernels below have four PRODUCER warps that expand (dp, argmax) straight
// into the shared-memory operand tile, in exactly the byte image TMA would have written (the tensor core
// applies the swizzle XOR on absolute shared-memory address bits when it reads), and
// publish it to the async proxy with fence.proxy.async + an mbarrier arrive per producer warp.
//
// ---- dgrad: "kw-scatter" implicit GEMM ----
// Measured on B200: an SS-mode 128 x N x 8 tf32 tcgen05.mma costs about (A bytes + B bytes) / 64 cycles, so
// small-N MMAs are shared-memory bound.  The three kw taps therefore sit on the N side:
//   D[pixel (y,x')][(kw,i)] = sum_kh sum_o dY[y-kh][x'][o] * W[kh][kw][o][i]          (N = 96)
//   da1[y][x][i]            = relu'(a1) * sum_kw D[(y, x-kw)][(kw,i)]
// one 128 x 96 x 8 MMA per (kh, k-step) re-uses the same A tile for all kw; the kh taps are accumulated by
// the tensor core through row-shifted views of the dY patch; the kw sum is two warp shuffles per value in
// the epilogue (a tile is 4 output rows x 32 pixels, a warp holds one row, lane = x).  Item = (image, half):
// half 0 -> output rows 0..15 (4 tiles), half 1 -> rows 16..25 (3 tiles); the patch holds dY rows
// y0-2 .. y0+15 x 32 pixels (24 real) for one 32-channel chunk, two chunks = a 2-deep pipeline.
// K-major operand, 128B swizzle: a pixel is one 128-byte row, and inside each 1024-byte group of 8 pixels
// the 16-byte chunk index is XORed with (pixel & 7):
//     byte offset of (row r, pixel x, channel quad cq) = r*4096 + x*128 + ((cq ^ (x & 7)) << 4)
// Columns 24..31 of every patch row are zero (written once at kernel start); rows outside the image are
// written as zeros for every item.
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
constexpr int DG4_UPT = (9 * WP * 8 + PROD_WARPS * 32 - 1) / (PROD_WARPS * 32);  // 7 work units per producer thread

// (registers capped at 104: 43 K per CTA leaves room for a co-resident CTA of the small reduction / communication
// kernels of the step, so those overlap this kernel; uncapped ptxas takes 128 x 416 = 53 K.  A cap of 88 costs this
// kernel 8 us -- 34.7 instead of 26.4 -- for nothing)
template <int MAXREG>
__global__ void __maxnreg__(MAXREG)
conv2_dgrad_tc_kernel(const __grid_constant__ CUtensorMap mapW, const float* __restrict__ dp,
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
  // Items alternate between 4-tile (upper half) and 3-tile (lower half) work.  With an even grid a CTA would
  // see the same half in every round (12 tiles against 9); swapping neighbours in odd rounds gives every CTA
  // both kinds (11 against 10).  nitems and the grid are even, so pairs stay inside a round.
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
        // work unit = (pooled row k of the patch, pooled column pw, channel quad cq): 9 x 12 x 8 = 864, up to 7
        // per thread; their global loads are issued before the patch slot is waited for
        float4 d[DG4_UPT];
        uchar4 a[DG4_UPT];
#pragma unroll
        for (int j = 0; j < DG4_UPT; ++j) {
          const int u = ptid + j * (PROD_WARPS * 32);
          const int cq = u & 7, pw = (u >> 3) % WP, k = u / (8 * WP), ph = ph0 + k;
          d[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          a[j] = make_uchar4(0, 0, 0, 0);
          if (u < 9 * WP * 8 && ph >= 0 && ph < HP) {
            const size_t e = (((size_t)n * HP + ph) * WP + pw) * C2 + kc * 32 + cq * 4;
            d[j] = *reinterpret_cast<const float4*>(dp + e);
            a[j] = *reinterpret_cast<const uchar4*>(amax + e);
          }
        }
        mbar_wait(&pempty[kc], (it & 1) ^ 1);
        uint8_t* pb = patch + kc * PATCH_BYTES;
#pragma unroll
        for (int j = 0; j < DG4_UPT; ++j) {
          const int u = ptid + j * (PROD_WARPS * 32);
          const int cq = u & 7, pw = (u >> 3) % WP, k = u / (8 * WP);
          if (u < 9 * WP * 8) {
#pragma unroll
            for (int s = 0; s < 4; ++s) {
              const int r = 2 * k + (s >> 1), x = 2 * pw + (s & 1);
              const float4 v = make_float4(a[j].x == s ? d[j].x : 0.f, a[j].y == s ? d[j].y : 0.f,
                                           a[j].z == s ? d[j].z : 0.f, a[j].w == s ? d[j].w : 0.f);
              *reinterpret_cast<float4*>(pb + r * ROW_BYTES + x * 128 + ((cq ^ (x & 7)) << 4)) = v;
            }
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
// wgrad:  dW[kh][kw][o][i] = sum_pixels dY[y][x][o] * a1[y+kh][x+kw][i]      (K = pixels)
// Item = (image, third) = 8 dY rows.  A operand = a1 rows y0..y0+9 by TMA (M = (kw, i): four 32-wide
// chunks one pixel apart, kw = 3 is a don't-care chunk), B operand = dY, N = 64 channels in two chunks,
// one k-step = 8 consecutive pixels of a row; three TMEM accumulators (kh) live for the whole kernel and
// a CTA leaves one partial row.  dY is never read from memory: four producer warps expand the pooled
// gradient dp[B][12][12][64] and the 2-bit arg-max codes straight into the MN-major operand tile (with
// the dgrad above, no kernel reads a dense dY any more).  MN-major tf32 tiles use the 128B swizzle
// with 32-byte atoms (tc_common.cuh): a pixel is one 128-byte row of 32 channels and the 32-byte chunk
// index is XORed with (pixel & 3):
//     byte offset of (row r, pixel x, channel c of the chunk) = r*4096 + x*128 + (((c >> 3) ^ (x & 3)) << 5) + (c & 7)*4
// Pixels 24..31 of a patch row are never an MMA operand (the K steps cover pixels 0..23), so they are
// left alone.  Over the whole grid the producers see every element of dp exactly once, which makes
// conv2's bias gradient (sum of dp over batch and pixels) a by-product: a thread always handles the same
// channel quad, keeps a running sum, and the CTA's total goes into the bias slots of its partial row.
// =============================================================================================
constexpr int WG_A_BYTES = 10 * P_PITCH * 128;  // 40,960
constexpr int WG_Y_BYTES = 8 * P_PITCH * 128;   // 32,768 per channel chunk
constexpr int WG_STAGE = WG_A_BYTES + 2 * WG_Y_BYTES;  // 106,496
constexpr int WG4_SMEM = 2 * WG_STAGE + 1024;
constexpr int CONV2_PART = 9 * C2 * C1 + C2;
constexpr int WG4_EPI = 4;
constexpr int WG4_THREADS = (WG4_EPI + PROD_WARPS + 1) * 32;  // 288
constexpr int WG4_UNITS = 4 * WP * 16;                         // (pooled row, pooled column, channel quad) = 768
constexpr int WG4_UPT = WG4_UNITS / (PROD_WARPS * 32);         // 6 units per producer thread

template <int MAXREG>
__global__ void __maxnreg__(MAXREG)  // register cap: see b200_conv2_dgrad_tc
conv2_wgrad_tc_kernel(const __grid_constant__ CUtensorMap mapA1, const float* __restrict__ dp,
                       const uint8_t* __restrict__ amax, float* __restrict__ part, int nitems) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2], tfull_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ float4 dbs[PROD_WARPS * 32];

  // warps 0..3 epilogue (id = TMEM sub-partition), 4..7 producers, 8 MMA issuer
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_epi = warp_id < WG4_EPI, is_mma = warp_id == WG4_EPI + PROD_WARPS;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA1);
    for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], 1 + PROD_WARPS); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tfull_bar, 1);
    fence_barrier_init();
  }
  if (is_mma) tmem_alloc(&tmem_base_smem, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (!is_epi && !is_mma) {
    const int ptid = threadIdx.x - WG4_EPI * 32;  // 0..127
    const int cq = ptid & 15;                      // channel quad 0..15 of this thread (128 % 16 == 0)
    const int kc = cq >> 3;                        // 32-channel chunk
    const uint32_t c32 = (uint32_t)(cq & 7) >> 1, cin = (uint32_t)(cq & 1) * 16u;  // 32-byte chunk, offset inside it
    const bool tma_warp = ptid < 32;               // warp-uniform
    const bool leader = tma_warp ? elect_one() : false;
    float4 dsum = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int it = 0; it < my_items; ++it) {
      const int item = blockIdx.x + it * gridDim.x, n = item / 3, third = item - n * 3, y0 = third * 8, s = it & 1;
      // the pooled gradient of this item goes to registers before the slot is waited for
      float4 d[WG4_UPT];
      uchar4 a[WG4_UPT];
#pragma unroll
      for (int j = 0; j < WG4_UPT; ++j) {
        const int u = ptid + j * (PROD_WARPS * 32), pw = (u >> 4) % WP, k = u / (16 * WP);
        const size_t e = (((size_t)n * HP + third * 4 + k) * WP + pw) * C2 + cq * 4;
        d[j] = *reinterpret_cast<const float4*>(dp + e);
        a[j] = *reinterpret_cast<const uchar4*>(amax + e);
      }
      mbar_wait(&empty_bar[s], ((it >> 1) & 1) ^ 1);
      uint8_t* st = smem + s * WG_STAGE;
      if (tma_warp) {
        if (leader) mbar_expect_tx(&full_bar[s], WG_A_BYTES);
        for (int r = 0; r < 10; ++r)
          if (leader) tma_load_4d(st + r * ROW_BYTES, &mapA1, &full_bar[s], 0, 0, y0 + r, n);
      }
      uint8_t* yb = st + WG_A_BYTES + kc * WG_Y_BYTES;
#pragma unroll
      for (int j = 0; j < WG4_UPT; ++j) {
        const int u = ptid + j * (PROD_WARPS * 32), pw = (u >> 4) % WP, k = u / (16 * WP);
        dsum.x += d[j].x; dsum.y += d[j].y; dsum.z += d[j].z; dsum.w += d[j].w;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = 2 * k + (q >> 1), x = 2 * pw + (q & 1);
          const float4 v = make_float4(a[j].x == q ? d[j].x : 0.f, a[j].y == q ? d[j].y : 0.f,
                                       a[j].z == q ? d[j].z : 0.f, a[j].w == q ? d[j].w : 0.f);
          *reinterpret_cast<float4*>(yb + r * ROW_BYTES + x * 128 + ((c32 ^ (uint32_t)(x & 3)) << 5) + cin) = v;
        }
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's operand reads
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_bar[s]);
    }
    // conv2 bias gradient of this CTA: 8 threads share a channel quad, combined in a fixed order
    dbs[ptid] = dsum;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (ptid < 16) {
      float4 t = dbs[ptid];
#pragma unroll
      for (int g = 1; g < 8; ++g) { const float4 v = dbs[ptid + 16 * g]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
      *reinterpret_cast<float4*>(part + (size_t)blockIdx.x * CONV2_PART + 9 * C2 * C1 + ptid * 4) = t;
    }
  } else if (is_mma) {
    const bool leader = elect_one();
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
  } else {
    const int q = warp_id;  // = kw of the rows this warp reads (q == 3: don't-care chunk)
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
  if (is_mma) tmem_dealloc(tmem_base, 256);
}

std::map<const void*, CUtensorMap> g_wmaps;
std::mutex g_mu;

// atom32 = 1: 32B-atom swizzle (N-major B operand of the dgrad); 0: plain 128B swizzle (K-major, forward)
int weight_map(const float* w, int atom32, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  const void* key = (const char*)w + atom32;  // weights are 256-float aligned: the low bit is free
  auto it = g_wmaps.find(key);
  if (it != g_wmaps.end()) { *out = it->second; return B200_OK; }
  if (g_wmaps.size() >= 64) g_wmaps.clear();  // bounded: a long-lived host may cycle through many buffers
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

// da1 = relu'(a1) * conv2_dgrad(max-pool-backward(dp, amax))
extern "C" int b200_conv2_dgrad_tc(const float* dp, const uint8_t* amax, const float* w, const float* a1, float* da1,
                                    int batch, void* stream) {
  B200_REQUIRE(batch > 0 && dp && amax && w && a1 && da1, "conv2_dgrad_tc: bad arguments");
  CUtensorMap mw;
  int r = weight_map(w, 1, &mw);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_tc_kernel<104>, cudaFuncAttributeMaxDynamicSharedMemorySize, DG4_SMEM));
    B200_CHECK(cudaFuncSetAttribute(conv2_dgrad_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, DG4_SMEM));
    attr = true;
  }
  const int nitems = 2 * batch;
  const int grid = min(nitems, b200_sm_count());
  // Register cap (B200_DGRAD_REGS = 104 | 128).  104 x 416 = 43 K registers would leave room for a co-resident CTA of
  // the small reduction / communication kernels; 128 is what ptxas takes when left alone (53 K: nothing fits beside
  // it).  Measured, us/step: 1 GPU 131.6 (104) vs 128.9 (128), 2 GPUs 146.2 vs 144.6 -- the kernel itself loses 4 us
  // to the cap and the overlap gives less back, so the default is 128.
  static const int regs = getenv("B200_DGRAD_REGS") ? atoi(getenv("B200_DGRAD_REGS")) : 128;
  if (regs >= 128)
    B200_CHECK(b200_launch(conv2_dgrad_tc_kernel<128>, dim3(grid), dim3(DG4_THREADS), (size_t)DG4_SMEM, (cudaStream_t)stream,
                           mw, dp, amax, a1, da1, nitems));
  else
    B200_CHECK(b200_launch(conv2_dgrad_tc_kernel<104>, dim3(grid), dim3(DG4_THREADS), (size_t)DG4_SMEM, (cudaStream_t)stream,
                           mw, dp, amax, a1, da1, nitems));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}


namespace {
struct A1Key {
  const void* p; int batch;
  bool operator<(const A1Key& o) const { return std::tie(p, batch) < std::tie(o.p, o.batch); }
};
std::map<A1Key, CUtensorMap> g_a1maps;

// a1 [B][26][26][32] as rows of 32 pixels (columns >= 26 zero-filled), 32B-atom swizzle (M-major A operand)
int a1_map(const float* a1, int batch, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  A1Key key{a1, batch};
  auto it = g_a1maps.find(key);
  if (it != g_a1maps.end()) { *out = it->second; return B200_OK; }
  if (g_a1maps.size() >= 64) g_a1maps.clear();  // bounded: a long-lived host may cycle through many buffers
  CUtensorMap m;
  uint64_t dims[4] = {32, 26, 26, (uint64_t)batch}, strides[3] = {128, 26 * 128, 676 * 128};
  uint32_t box[4] = {32, P_PITCH, 1, 1};
  int r = tcmap::encode(&m, a1, 4, dims, strides, box, 1);
  if (r != B200_OK) return r;
  g_a1maps[key] = m;
  *out = m;
  return B200_OK;
}
}  // namespace

// conv2 weight + bias gradient straight from the pooled gradient (no dense dY tensor):
//   part[grid][18496]: per-CTA partial sums, dW[tap][o][i] in floats [0, 18432), db[o] in [18432, 18496);
//   grid = b200_conv2_wgrad_tc_parts(batch).  dw / db (optional): fixed-order reduction of the partials.
extern "C" int b200_conv2_wgrad_tc_parts(int batch) { return min(3 * batch, b200_sm_count()); }
extern "C" int b200_conv2_wgrad_tc(const float* dp, const uint8_t* amax, const float* a1, float* part, float* dw,
                                    float* db, int batch, void* stream) {
  B200_REQUIRE(batch > 0 && dp && amax && a1 && part, "conv2_wgrad_tc: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  CUtensorMap ma;
  int r = a1_map(a1, batch, &ma);
  if (r != B200_OK) return r;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(conv2_wgrad_tc_kernel<80>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG4_SMEM));
    B200_CHECK(cudaFuncSetAttribute(conv2_wgrad_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG4_SMEM));
    attr = true;
  }
  const int nitems = 3 * batch;
  const int grid = min(nitems, b200_sm_count());
  static const int regs = getenv("B200_WGRAD_REGS") ? atoi(getenv("B200_WGRAD_REGS")) : 128;  // 80 | 128; measured 129.4 vs 128.9 us/step
  if (regs >= 128)
    B200_CHECK(b200_launch(conv2_wgrad_tc_kernel<128>, dim3(grid), dim3(WG4_THREADS), (size_t)WG4_SMEM, s, ma, dp, amax, part,
                           nitems));
  else
    B200_CHECK(b200_launch(conv2_wgrad_tc_kernel<80>, dim3(grid), dim3(WG4_THREADS), (size_t)WG4_SMEM, s, ma, dp, amax, part,
                           nitems));
  b200_count_launch();
  if (dw) b200_launch_reduce_partials(dw, part, 9 * C2 * C1, grid, CONV2_PART, s);
  if (db) b200_launch_reduce_partials(db, part + 9 * C2 * C1, C2, grid, CONV2_PART, s);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
