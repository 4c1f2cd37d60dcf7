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

// conv2's half of the conv bucket on one GPU: reduce the wgrad kernel's per-CTA partial rows into the
// gradient buffer.  Runs on the wgrad branch, under conv2 dgrad / conv1 wgrad (11 MB of partials at 148
// CTAs would otherwise be read by the step's last kernel).  One CTA per FIN_CHUNK-float chunk.
__global__ void __launch_bounds__(FIN_THREADS)
reduce_segs_kernel(float* __restrict__ dst, FinalizeSegs segs) {
  __shared__ float4 red[FIN_SLICES][FIN_COLS];
  pdl_trigger();
  pdl_wait();
  const int off = blockIdx.x * FIN_CHUNK;
  const float4 t = reduce_partial_chunk(segs, off, red);
  if (threadIdx.x < FIN_COLS) *reinterpret_cast<float4*>(dst + off + 4 * threadIdx.x) = t;
}

// Single-GPU end-of-step kernel.  CTAs [0, nchunks) each own FIN_CHUNK floats of the conv bucket: chunks below
// first_partial already have their gradient in g (conv2, reduced on the wgrad branch), the others (conv1)
// reduce their partial rows in a fixed order first; then Adadelta.  The remaining CTAs stream over the fc
// bucket unless that was updated on the fc branch already.  The last CTA to finish advances the step.
__global__ void __launch_bounds__(FIN_THREADS)
adadelta_finalize_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ v, float* __restrict__ u,
                         int b0_n4, int b1_off, int nchunks, int first_partial, FinalizeSegs segs,
                         B200Ctrl* __restrict__ ctrl, float rho, float eps, int advance,
                         unsigned int* __restrict__ done_counter) {
  __shared__ float4 red[FIN_SLICES][FIN_COLS];
  pdl_trigger();
  pdl_wait();
  const float lr = ctrl->lr;
  if ((int)blockIdx.x < nchunks) {
    const int col = threadIdx.x % FIN_COLS, off = blockIdx.x * FIN_CHUNK;
    const size_t i = (size_t)b1_off + off + 4 * col;
    float4 t;
    if ((int)blockIdx.x >= first_partial) {
      t = reduce_partial_chunk(segs, off, red);
      if (threadIdx.x < FIN_COLS) *reinterpret_cast<float4*>(g + i) = t;  // keep the gradient buffer complete (tests, debugging)
    } else if (threadIdx.x < FIN_COLS) {
      t = *reinterpret_cast<const float4*>(g + i);
    }
    if (threadIdx.x < FIN_COLS) {
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

int b200_reduce_segs(float* dst, int nchunks, const FinalizeSegs& segs, cudaStream_t s) {
  B200_REQUIRE(dst && nchunks > 0, "reduce_segs: bad arguments");
  B200_CHECK(b200_launch(reduce_segs_kernel, dim3(nchunks), dim3(FIN_THREADS), 0, s, dst, segs));
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// bucket 0 = floats [0, b1_off) (fc), bucket 1 = floats [b1_off, b1_off + FIN_CHUNK*nchunks) (conv)
int b200_adadelta_finalize(float* p, float* g, float* v, float* u, long b1_off, int nchunks, int first_partial,
                           const FinalizeSegs& segs, void* ctrl, float rho, float eps, int advance,
                           unsigned int* done_counter, bool bucket0_done, cudaStream_t s) {
  B200_REQUIRE(b1_off % 4 == 0 && nchunks > 0 && first_partial >= 0 && first_partial <= nchunks,
               "adadelta_finalize: bad bucket layout");
  // bucket0_done: the fc bucket was already updated elsewhere (engine branch stream); only the conv chunks run
  const int grid = nchunks + (bucket0_done ? 0 : 2 * b200_sm_count());
  B200_CHECK(b200_launch(adadelta_finalize_kernel, dim3(grid), dim3(FIN_THREADS), 0, s, p, g, v, u,
                         bucket0_done ? 0 : (int)(b1_off / 4), (int)b1_off, nchunks, first_partial, segs,
                         (B200Ctrl*)ctrl, rho, eps, advance, done_counter));
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
