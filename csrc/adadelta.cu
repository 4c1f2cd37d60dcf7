// Single-GPU fused Adadelta over the flat parameter buffer.
//
// torch.optim.Adadelta on CUDA runs ~12 multi-tensor kernels per step (mnist_ddp.py:73,176);
// this is one pass: read g, v, u, p — write v, u, p (28 B/element).  The multi-GPU variant lives
// in allreduce.cu, fused with the cross-GPU reduction.  The update follows
// LIB:torch/optim/adadelta.py:292-302 (rho = .9, eps = 1e-6, weight_decay = 0):
//     v   = rho v + (1-rho) g^2
//     d   = sqrt(u + eps) / sqrt(v + eps) * g
//     u   = rho u + (1-rho) d^2
//     p  -= lr d
// The last block to finish also advances the control block (step counter, sample cursor), which
// makes this kernel the "end of step" marker inside the replayed CUDA graph.
#include "common.cuh"
#include "engine.h"

__global__ void __launch_bounds__(256)
adadelta_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ v, float* __restrict__ u,
                int n4, B200Ctrl* __restrict__ ctrl, float rho, float eps, int advance_batch,
                unsigned int* __restrict__ done_counter) {
  pdl_trigger();
  pdl_wait();
  const float lr = ctrl->lr;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n4; e += gridDim.x * blockDim.x) {
    const float4 gg = reinterpret_cast<const float4*>(g)[e];
    float4 pp = reinterpret_cast<float4*>(p)[e], vv = reinterpret_cast<float4*>(v)[e],
           uu = reinterpret_cast<float4*>(u)[e];
    adadelta_update(gg.x, pp.x, vv.x, uu.x, lr, rho, eps);
    adadelta_update(gg.y, pp.y, vv.y, uu.y, lr, rho, eps);
    adadelta_update(gg.z, pp.z, vv.z, uu.z, lr, rho, eps);
    adadelta_update(gg.w, pp.w, vv.w, uu.w, lr, rho, eps);
    reinterpret_cast<float4*>(p)[e] = pp;
    reinterpret_cast<float4*>(v)[e] = vv;
    reinterpret_cast<float4*>(u)[e] = uu;
  }
  if (advance_batch >= 0 && done_counter) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(done_counter, 1u) == gridDim.x - 1) {  // every block has read ctrl->lr by now
        *done_counter = 0;
        ctrl->step += 1;
        ctrl->pos += advance_batch;
      }
    }
  }
}

// Single-GPU end-of-step kernel when the conv-bucket gradients are still per-CTA partial sums:
// CTAs [0, nchunks) each own 256 floats of the conv bucket (reduce partials in a fixed order, then
// Adadelta); the remaining CTAs stream over the fc bucket.  Replaces four reduction launches plus the
// plain Adadelta launch.
__global__ void __launch_bounds__(1024)
adadelta_finalize_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ v, float* __restrict__ u,
                         int b0_n4, int b1_off, int nchunks, FinalizeSegs segs, B200Ctrl* __restrict__ ctrl, float rho,
                         float eps, int advance, unsigned int* __restrict__ done_counter) {
  __shared__ float4 red[FIN_SLICES][64];
  pdl_trigger();
  pdl_wait();
  const float lr = ctrl->lr;
  if ((int)blockIdx.x < nchunks) {
    const int col = threadIdx.x & 63, slice = threadIdx.x >> 6, off = blockIdx.x * 256;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const FinalizeSegs::Seg sg = segs.seg[k];
      const int e = off - sg.dst_off + 4 * col;
      if (sg.src && e >= 0 && e < sg.n) {
#pragma unroll 8
        for (int q = slice; q < sg.nparts; q += FIN_SLICES) {
          const float4 x = *reinterpret_cast<const float4*>(sg.src + (size_t)q * sg.stride + e);
          acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
        }
      }
    }
    red[slice][col] = acc;
    __syncthreads();
    if (slice == 0) {
      float4 t = red[0][col];
#pragma unroll
      for (int q = 1; q < FIN_SLICES; ++q) { const float4 x = red[q][col]; t.x += x.x; t.y += x.y; t.z += x.z; t.w += x.w; }
      const size_t i = (size_t)b1_off + off + 4 * col;
      *reinterpret_cast<float4*>(g + i) = t;  // keep the gradient buffer complete (tests, debugging)
      float4 pp = *reinterpret_cast<float4*>(p + i), vv = *reinterpret_cast<float4*>(v + i),
             uu = *reinterpret_cast<float4*>(u + i);
      adadelta_update(t.x, pp.x, vv.x, uu.x, lr, rho, eps);
      adadelta_update(t.y, pp.y, vv.y, uu.y, lr, rho, eps);
      adadelta_update(t.z, pp.z, vv.z, uu.z, lr, rho, eps);
      adadelta_update(t.w, pp.w, vv.w, uu.w, lr, rho, eps);
      *reinterpret_cast<float4*>(p + i) = pp;
      *reinterpret_cast<float4*>(v + i) = vv;
      *reinterpret_cast<float4*>(u + i) = uu;
    }
  } else {
    const int nb = gridDim.x - nchunks;
    for (int e = (blockIdx.x - nchunks) * blockDim.x + threadIdx.x; e < b0_n4; e += nb * blockDim.x) {
      const float4 gg = reinterpret_cast<const float4*>(g)[e];
      float4 pp = reinterpret_cast<float4*>(p)[e], vv = reinterpret_cast<float4*>(v)[e],
             uu = reinterpret_cast<float4*>(u)[e];
      adadelta_update(gg.x, pp.x, vv.x, uu.x, lr, rho, eps);
      adadelta_update(gg.y, pp.y, vv.y, uu.y, lr, rho, eps);
      adadelta_update(gg.z, pp.z, vv.z, uu.z, lr, rho, eps);
      adadelta_update(gg.w, pp.w, vv.w, uu.w, lr, rho, eps);
      reinterpret_cast<float4*>(p)[e] = pp;
      reinterpret_cast<float4*>(v)[e] = vv;
      reinterpret_cast<float4*>(u)[e] = uu;
    }
  }
  if (advance >= 0) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(done_counter, 1u) == gridDim.x - 1) {
        *done_counter = 0;
        ctrl->step += 1;
        ctrl->pos += advance;
      }
    }
  }
}

// bucket 0 = floats [0, b1_off) (fc), bucket 1 = floats [b1_off, b1_off + 256*nchunks) (conv, as partials)
int b200_adadelta_finalize(float* p, float* g, float* v, float* u, long b1_off, int nchunks, const FinalizeSegs& segs,
                           void* ctrl, float rho, float eps, int advance, unsigned int* done_counter,
                           bool bucket0_done, cudaStream_t s) {
  B200_REQUIRE(b1_off % 4 == 0 && nchunks > 0, "adadelta_finalize: bad bucket layout");
  // bucket0_done: the fc bucket was already updated elsewhere (engine branch stream); only the conv chunks run
  const int grid = nchunks + (bucket0_done ? 0 : 2 * b200_sm_count());
  B200_CHECK(b200_launch(adadelta_finalize_kernel, dim3(grid), dim3(64 * FIN_SLICES), 0, s, p, g, v, u,
                         bucket0_done ? 0 : (int)(b1_off / 4),
                         (int)b1_off, nchunks, segs, (B200Ctrl*)ctrl, rho, eps, advance, done_counter));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// n must be a multiple of 4 (the flat layout pads every segment to 256 floats).
// advance_batch >= 0: also advance ctrl (step += 1, pos += advance_batch); needs done_counter.
extern "C" int b200_adadelta(float* p, const float* g, float* v, float* u, long n, void* ctrl, float rho, float eps,
                             int advance_batch, unsigned int* done_counter, void* stream) {
  B200_REQUIRE(n > 0 && n % 4 == 0, "adadelta: n (%ld) must be a positive multiple of 4", n);
  const int n4 = (int)(n / 4);
  const int grid = min(ceil_div(n4, 256), 4 * b200_sm_count());
  B200_CHECK(b200_launch(adadelta_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, p, g, v, u, n4, (B200Ctrl*)ctrl,
                         rho, eps, advance_batch, done_counter));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
