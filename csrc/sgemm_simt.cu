// Generic fp32 SIMT GEMM with arbitrary operand strides: C[m][n] = sum_k A(m,k) * B(k,n).
//
// Used as the fp32-exact path for the fc1 layer (the reference runs fc1/fc2 in full fp32,
// SURVEY H3) and as the oracle for the tcgen05 TF32 GEMM in tc_gemm.cu:
//   fc1 forward : A = act[B][9216] (k contiguous), B(k,n) = W1[n][k]         -> split-K partials
//   fc1 dgrad   : A = dh[B][128],  B(k,n) = W1[k][n], epilogue masks by the forward activation
//   fc1 wgrad   : A(m,k) = dh[k][m], B(k,n) = act[k][n]
// 64x64x16 tiles, 256 threads, 4x4 outputs per thread, split-K over gridDim.z.
#include "common.cuh"

constexpr int BM = 64, BN = 64, BK = 16;

// Load a [ROWS x BK] tile where element (r, k) lives at base[r*sr + k*sk].  The thread mapping
// puts the unit-stride dimension across consecutive threads so global loads coalesce.
template <bool K_CONTIG>
__device__ __forceinline__ void load_tile(float (*dst)[BM + 4], const float* __restrict__ base, long sr, long sk,
                                          int r0, int k0, int rmax, int kmax) {
#pragma unroll
  for (int it = 0; it < (BM * BK) / 256; ++it) {
    const int e = it * 256 + threadIdx.x;
    int r, k;
    if (K_CONTIG) { k = e % BK; r = e / BK; } else { r = e % BM; k = e / BM; }
    const int gr = r0 + r, gk = k0 + k;
    dst[k][r] = (gr < rmax && gk < kmax) ? base[gr * sr + gk * sk] : 0.f;
  }
}

// EPI: 0 = store, 1 = store acc * (aux[m][n] > 0 ? scale : 0)
template <bool A_KC, bool B_KC, int EPI>
__global__ void __launch_bounds__(256)
sgemm_kernel(const float* __restrict__ A, long sam, long sak, const float* __restrict__ B, long sbk, long sbn,
             float* __restrict__ C, int M, int N, int K, int kchunk, const float* __restrict__ aux, float scale) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int kbeg = blockIdx.z * kchunk, kend = min(K, kbeg + kchunk);
  const int tm = (threadIdx.x >> 4) * 4, tn = (threadIdx.x & 15) * 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = kbeg; k0 < kend; k0 += BK) {
    load_tile<A_KC>(As, A, sam, sak, m0, k0, M, kend);
    load_tile<B_KC>(Bs, B, sbn, sbk, n0, k0, N, kend);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][tm]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tn]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* out = C + (size_t)blockIdx.z * M * N;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + tm + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tn + j;
      if (n >= N) continue;
      float v = acc[i][j];
      if (EPI == 1) v = aux[(size_t)m * N + n] > 0.f ? v * scale : 0.f;
      out[(size_t)m * N + n] = v;
    }
  }
}

// a_kc / b_kc: whether k is the unit-stride dimension of A / B.  splits > 1 writes `splits`
// partial C matrices back to back (caller reduces).  epi = 1 needs aux (same shape as C).
extern "C" int b200_sgemm(const float* A, long sam, long sak, const float* B, long sbk, long sbn, float* C,
                          int M, int N, int K, int splits, int epi, const float* aux, float scale, void* stream) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0 && splits > 0, "sgemm: bad shape M=%d N=%d K=%d splits=%d", M, N, K, splits);
  B200_REQUIRE(epi == 0 || (epi == 1 && aux && splits == 1), "sgemm: epilogue %d needs aux and splits == 1", epi);
  const bool akc = (sak == 1), bkc = (sbk == 1);
  int kchunk = ceil_div(ceil_div(K, splits), BK) * BK;
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), splits);
  cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH(AK, BKC, E) sgemm_kernel<AK, BKC, E><<<grid, 256, 0, s>>>(A, sam, sak, B, sbk, sbn, C, M, N, K, kchunk, aux, scale)
  if (epi == 0) {
    if (akc && bkc) LAUNCH(true, true, 0);
    else if (akc) LAUNCH(true, false, 0);
    else if (bkc) LAUNCH(false, true, 0);
    else LAUNCH(false, false, 0);
  } else {
    if (akc && bkc) LAUNCH(true, true, 1);
    else if (akc) LAUNCH(true, false, 1);
    else if (bkc) LAUNCH(false, true, 1);
    else LAUNCH(false, false, 1);
  }
#undef LAUNCH
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
