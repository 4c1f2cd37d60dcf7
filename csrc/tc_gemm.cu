// TF32 GEMM on the 5th-generation tensor cores: TMA -> shared memory -> tcgen05.mma -> TMEM ->
// tcgen05.ld epilogue.  D[M][N] = A . B with fp32 accumulation; either operand may be K-major or
// MN-major, which is what lets the SAME kernel serve the three fc1 GEMMs without transposed copies:
//     fc1 forward  D[B][128]    = act[B][9216] . W1[128][9216]^T   (K-major, K-major, split-K)
//     fc1 wgrad    D[128][9216] = dh[B][128]^T . act[B][9216]       (M-major, N-major, K = batch)
//     fc1 dgrad    D[B][9216]   = dh[B][128]   . W1[128][9216]      (K-major, N-major) + mask epilogue
// (mnist_ddp.py:57 and its backward).
//
// One CTA computes one 128 x BN tile over its K range: one TMA producer warp, one MMA issuer warp (one
// elected lane; TMEM owner) and 8 epilogue warps (two per TMEM sub-partition).  An mbarrier ring of 32-wide
// K blocks decouples TMA from the tensor core.
//
// SPLIT = true is the fp32-class variant ("3xTF32"): the reference runs fc1 as an fp32 SGEMM
// (torch.backends.cuda.matmul.allow_tf32 = False), and a single TF32 pass keeps only 10 mantissa
// bits of each operand.  Every operand x is split into hi = x with its 13 low mantissa bits cleared (what the
// tensor core reads anyway: it truncates) and lo = x - hi (exact in fp32, |lo| < 2^-10 |x|), and the product is
// accumulated as hi*hi + hi*lo + lo*hi into the same TMEM accumulator; the dropped terms are O(2^-20) relative.
// The split is done on chip: while the mainloop runs, the (otherwise idle) epilogue warps read each TMA-written
// stage and write lo to a second tile at the same swizzled offsets, then hand the stage to the MMA warp
// (fence.proxy.async + mbarrier).  Nothing extra is read from or written to global memory.  What remains is the
// tensor core's own accumulator, which truncates when it adds: on long K chains (thousands of accumulating MMAs
// into one output) that error reaches a few 1e-5 relative; the engine keeps chains short (split-K, K = 128, K = batch).
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "tc_common.cuh"

using namespace tc;

namespace tcmap {
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeFn get_encode() {
  static EncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeFn)p;
  }
  return fn;
}

int encode(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
           const uint32_t* box, int swizzle) {
  EncodeFn fn = get_encode();
  if (!fn) {
    b200_set_error("cuTensorMapEncodeTiled entry point not available");
    return B200_ERR_UNSUPPORTED;
  }
  cuuint64_t gd[5], gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200_set_error("cuTensorMapEncodeTiled failed (CUresult %d), rank %d dims %llu %llu %llu box %u %u %u", (int)r, rank,
                   (unsigned long long)gd[0], (unsigned long long)gd[1], (unsigned long long)(rank > 2 ? gd[2] : 0), bx[0],
                   bx[1], rank > 2 ? bx[2] : 0);
    return B200_ERR_CUDA;
  }
  return B200_OK;
}
}  // namespace tcmap

namespace {

constexpr int BM = 128, BK = 32;
constexpr int A_BYTES = BM * BK * 4;  // 16 KiB
__host__ __device__ constexpr int stages_for(bool split, int bn) { return split && bn > 64 ? 3 : 4; }  // <= 193 KiB

template <int BN, bool A_MN, bool B_MN, int EPI, bool SPLIT>
__global__ void __launch_bounds__(320, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
               float* __restrict__ D, int M, int N, int nkb_total, int kb_per_split, const float* __restrict__ aux,
               float scale) {
  constexpr int STAGES = stages_for(SPLIT, BN);
  constexpr int B_BYTES = BN * BK * 4;
  constexpr int HI_BYTES = A_BYTES + B_BYTES;                 // what TMA writes per stage
  constexpr int STAGE_BYTES = SPLIT ? 2 * HI_BYTES : HI_BYTES;  // SPLIT: [A hi | B hi | A lo | B lo]
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);  // SW128 needs 1 KiB alignment
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], split_bar[STAGES], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;

  // Roles by warp: the TMA producer and the MMA issuer get the two HIGHEST warp ids of the CTA (the warp
  // arbiter favours high ids, and each then owns an SM sub-partition's top priority); the epilogue warps are
  // ids 0.. so that id % 4 is their TMEM sub-partition.  `warp` below is the role index (0 producer, 1 MMA,
  // 2.. epilogue), `wq` the TMEM sub-partition of an epilogue warp.
  const int warp_id = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, wq = warp_id & 3;
  const int warp = warp_id == nwarps - 2 ? 0 : (warp_id == nwarps - 1 ? 1 : warp_id + 2);
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
  const int kb0 = blockIdx.z * kb_per_split;
  const int nkb = min(kb_per_split, nkb_total - kb0);

  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&mapA);
    tma_prefetch_desc(&mapB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&split_bar[s], nwarps - 2);
    }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();  // everything above is on-chip set-up; global memory is only touched from here on

  if (warp == 0) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES;
        mbar_wait(&empty_bar[s], ((kb / STAGES) & 1) ^ 1);
        uint8_t* sa = smem + s * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
        if (leader) mbar_expect_tx(&full_bar[s], HI_BYTES);
        const int k0 = (kb0 + kb) * BK;
        if (leader) {
          if (A_MN) tma_load_3d(sa, &mapA, &full_bar[s], 0, k0, m0 / 32);
          else tma_load_2d(sa, &mapA, &full_bar[s], k0, m0);
          if (B_MN) tma_load_3d(sb, &mapB, &full_bar[s], 0, k0, n0 / 32);
          else tma_load_2d(sb, &mapB, &full_bar[s], k0, n0);
        }
      }
    }
  } else if (warp == 1) {
    {
      const bool leader = elect_one();  // whole warp stays convergent: operands live in uniform registers
      constexpr uint32_t idesc = idesc_tf32(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES;
        mbar_wait(SPLIT ? &split_bar[s] : &full_bar[s], (kb / STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * STAGE_BYTES), sb = sa + A_BYTES;
#pragma unroll
        for (int kk = 0; kk < BK / 8; ++kk) {
          const uint64_t ad = A_MN ? desc_mnmajor(sa + kk * 1024, BK * 128) : desc_kmajor(sa + kk * 32);
          const uint64_t bd = B_MN ? desc_mnmajor(sb + kk * 1024, BK * 128) : desc_kmajor(sb + kk * 32);
          if (leader) umma_tf32(tmem_base, ad, bd, idesc, (kb | kk) != 0);
          if (SPLIT) {  // + hi.lo + lo.hi: the lo tiles sit HI_BYTES further, same layout
            const uint64_t al = A_MN ? desc_mnmajor(sa + HI_BYTES + kk * 1024, BK * 128) : desc_kmajor(sa + HI_BYTES + kk * 32);
            const uint64_t bl = B_MN ? desc_mnmajor(sb + HI_BYTES + kk * 1024, BK * 128) : desc_kmajor(sb + HI_BYTES + kk * 32);
            if (leader) umma_tf32(tmem_base, ad, bl, idesc, 1u);
            if (leader) umma_tf32(tmem_base, al, bd, idesc, 1u);
          }
        }
        if (leader) umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs have read it
      }
      if (leader) umma_commit(&tmem_full_bar);
    }
  } else {
    // 8 epilogue warps (or 4): warps w and w+4 share TMEM sub-partition q and take alternate 32-column
    // chunks of the tile
    const int q = wq;  // TMEM sub-partition this warp may read
    const int h = (warp - 2) >> 2, cstep = 32 * ((nwarps - 2) >> 2);
    const int m = m0 + q * 32 + lane;
    float* drow = D + ((size_t)blockIdx.z * M + m) * N;
    const float* arow = EPI == 1 ? aux + (size_t)m * N : nullptr;
    // EPI 1: the mask operand does not depend on the MMAs; its (row-strided, slow) loads are issued one
    // chunk ahead, the first of them before the accumulator is waited for
    float4 anext[8];
    auto load_aux = [&](int c0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = n0 + c0 + 4 * j;
        anext[j] = (m < M && n < N) ? *reinterpret_cast<const float4*>(arow + n) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    if (EPI == 1) load_aux(h * 32);
    if (SPLIT) {
      // operand split, stage by stage, by all epilogue threads: lo = x - (x & ~0x1fff).  The hi tile is the raw tile:
      // kind::tf32 reads an fp32 operand by TRUNCATING it to TF32 (measured, tests/test_gpu_kernels.py:
      // test_tf32_operand_rounding_probe), so the tensor core itself sees x & ~0x1fff there.
      const int et = (warp - 2) * 32 + lane, nthr = (nwarps - 2) * 32;
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES;
        mbar_wait(&full_bar[s], (kb / STAGES) & 1);
        uint8_t* st = smem + s * STAGE_BYTES;
#pragma unroll 4
        for (int o = et * 16; o < HI_BYTES; o += nthr * 16) {
          const uint4 x = *reinterpret_cast<const uint4*>(st + o);
          const uint4 hi = make_uint4(x.x & 0xffffe000u, x.y & 0xffffe000u, x.z & 0xffffe000u, x.w & 0xffffe000u);
          const float4 lo = make_float4(__uint_as_float(x.x) - __uint_as_float(hi.x), __uint_as_float(x.y) - __uint_as_float(hi.y),
                                        __uint_as_float(x.z) - __uint_as_float(hi.z), __uint_as_float(x.w) - __uint_as_float(hi.w));
          *reinterpret_cast<float4*>(st + HI_BYTES + o) = lo;
        }
        fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's operand reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&split_bar[s]);
      }
    }
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c0 = h * 32; c0 < BN; c0 += cstep) {
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + c0, v);
      float4 acur[8];
      if (EPI == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acur[j] = anext[j];
        if (c0 + cstep < BN) load_aux(c0 + cstep);
      }
      if (m < M) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int n = n0 + c0 + j;
          if (n < N) {
            float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            if (EPI == 1) {
              const float4 a = acur[j >> 2];
              o.x = a.x > 0.f ? o.x * scale : 0.f; o.y = a.y > 0.f ? o.y * scale : 0.f;
              o.z = a.z > 0.f ? o.z * scale : 0.f; o.w = a.w > 0.f ? o.w * scale : 0.f;
            }
            *reinterpret_cast<float4*>(drow + n) = o;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN);
}

struct MapKey {
  const void* p; int rows, cols, mn, tile;
  bool operator<(const MapKey& o) const {
    return std::tie(p, rows, cols, mn, tile) < std::tie(o.p, o.rows, o.cols, o.mn, o.tile);
  }
};
std::map<MapKey, CUtensorMap> g_maps;
std::mutex g_maps_mu;

// Operand with logical shape [R][K] (R = M or N).  mn = 0: stored [R][K]; mn = 1: stored [K][R].
int get_map(const float* p, int R, int K, int mn, int tile, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_maps_mu);
  MapKey key{p, R, K, mn, tile};
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return B200_OK; }
  if (g_maps.size() >= 256) g_maps.clear();  // bounded: a long-lived host may cycle through many buffers
  CUtensorMap m;
  int r;
  if (!mn) {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)R}, strides[1] = {(uint64_t)K * 4};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)tile};
    r = tcmap::encode(&m, p, 2, dims, strides, box);
  } else {
    uint64_t dims[3] = {32, (uint64_t)K, (uint64_t)(R / 32)}, strides[2] = {(uint64_t)R * 4, 128};
    uint32_t box[3] = {32, (uint32_t)BK, (uint32_t)(tile / 32)};
    r = tcmap::encode(&m, p, 3, dims, strides, box, 1);
  }
  if (r != B200_OK) return r;
  g_maps[key] = m;
  *out = m;
  return B200_OK;
}

template <int BN, bool A_MN, bool B_MN, int EPI, bool SPLIT>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, float* D, int M, int N, int nkb, int kbs, int splits,
           const float* aux, float scale, cudaStream_t s) {
  constexpr int smem = stages_for(SPLIT, BN) * (SPLIT ? 2 : 1) * (A_BYTES + BN * BK * 4) + 1024;
  static bool attr = false;
  if (!attr) {
    B200_CHECK(cudaFuncSetAttribute(tc_gemm_kernel<BN, A_MN, B_MN, EPI, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr = true;
  }
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), splits);
  static const int ew = getenv("B200_GEMM_EW") ? atoi(getenv("B200_GEMM_EW")) : 8;  // epilogue warps: 4 or 8
  B200_CHECK(b200_launch(tc_gemm_kernel<BN, A_MN, B_MN, EPI, SPLIT>, grid, dim3(ew == 4 ? 192 : 320), (size_t)smem, s, ma, mb, D,
                         M, N, nkb, kbs, aux, scale));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <int BN, int EPI, bool SPLIT>
int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ma, const CUtensorMap& mb, float* D, int M, int N, int nkb,
                   int kbs, int splits, const float* aux, float scale, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<BN, false, false, EPI, SPLIT>(ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
  if (!a_mn && b_mn) return launch<BN, false, true, EPI, SPLIT>(ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
  if (a_mn && b_mn) return launch<BN, true, true, EPI, SPLIT>(ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
  return launch<BN, true, false, EPI, SPLIT>(ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
}

template <bool SPLIT>
int dispatch_tile(int bn, int epi, int a_mn, int b_mn, const CUtensorMap& ma, const CUtensorMap& mb, float* D, int M,
                  int N, int nkb, int kbs, int splits, const float* aux, float scale, cudaStream_t s) {
  if (bn == 128) {
    return epi ? dispatch_major<128, 1, SPLIT>(a_mn, b_mn, ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s)
               : dispatch_major<128, 0, SPLIT>(a_mn, b_mn, ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
  }
  return epi ? dispatch_major<64, 1, SPLIT>(a_mn, b_mn, ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s)
             : dispatch_major<64, 0, SPLIT>(a_mn, b_mn, ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
}

}  // namespace

extern "C" int b200_has_tc(void) { return 1; }

// fp32_class != 0: every operand is split into TF32 hi + lo parts on chip and the product accumulated as
// hi.hi + hi.lo + lo.hi (see the file header); 0: one TF32 pass
extern "C" int b200_tc_gemm_ex(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn,
                               int splits, int epi, const float* aux, float scale, int fp32_class, void* stream) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0 && splits > 0, "tc_gemm: bad shape M=%d N=%d K=%d splits=%d", M, N, K, splits);
  B200_REQUIRE(N % 4 == 0, "tc_gemm: N (%d) must be a multiple of 4", N);
  B200_REQUIRE(epi == 0 || (epi == 1 && aux && splits == 1), "tc_gemm: epilogue %d needs aux and splits == 1", epi);
  B200_REQUIRE(a_mn ? (M % 32 == 0) : (K % 4 == 0), "tc_gemm: A layout needs %s", a_mn ? "M % 32 == 0" : "K % 4 == 0");
  B200_REQUIRE(b_mn ? (N % 32 == 0) : (K % 4 == 0), "tc_gemm: B layout needs %s", b_mn ? "N % 32 == 0" : "K % 4 == 0");
  B200_REQUIRE((((uintptr_t)A | (uintptr_t)B | (uintptr_t)D) & 15) == 0, "tc_gemm: pointers must be 16-byte aligned");
  const int nkb = ceil_div(K, BK);
  B200_REQUIRE(splits <= nkb, "tc_gemm: %d splits but only %d K blocks of %d", splits, nkb, BK);
  int kbs = ceil_div(nkb, splits);
  B200_REQUIRE((splits - 1) * kbs < nkb, "tc_gemm: %d splits leave an empty K range (%d K blocks)", splits, nkb);
  const int bn = (ceil_div(M, BM) * ceil_div(N, 128) * splits >= 120) ? 128 : 64;
  CUtensorMap ma, mb;
  int r = get_map(A, M, K, a_mn, BM, &ma);
  if (r != B200_OK) return r;
  r = get_map(B, N, K, b_mn, bn, &mb);
  if (r != B200_OK) return r;
  cudaStream_t s = (cudaStream_t)stream;
  return fp32_class ? dispatch_tile<true>(bn, epi, a_mn, b_mn, ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s)
                    : dispatch_tile<false>(bn, epi, a_mn, b_mn, ma, mb, D, M, N, nkb, kbs, splits, aux, scale, s);
}

extern "C" int b200_tc_gemm(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn,
                            int splits, int epi, const float* aux, float scale, void* stream) {
  return b200_tc_gemm_ex(A, B, D, M, N, K, a_mn, b_mn, splits, epi, aux, scale, 0, stream);
}
