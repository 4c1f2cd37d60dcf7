// Fused gradient all-reduce + 1/world averaging + Adadelta over NVLink peer memory.
//
// This is the replacement for DistributedDataParallel's communication path (mnist_ddp.py:173 and
// what fires inside loss.backward() / optimizer.step(), mnist_ddp.py:72-73; SURVEY N15, N17, C1-C4,
// K25, section 8 rows 11, 12, 14, 17).  Per gradient bucket the reference runs: 4 bucket-copy kernels
// with the 1/W scale, one ncclAllReduce, 4 copy-back kernels, and later ~12 multi-tensor optimizer
// kernels.  Here a bucket is ONE kernel: every CTA joins a cross-GPU barrier, then loads the W ranks'
// gradient slices straight from their (IPC-mapped) buckets, sums them in a fixed rank order — so all
// ranks get bit-identical results — scales, applies the Adadelta update and writes parameters.
//
//   one-shot : every rank reduces the whole bucket and updates its full replica.  (W-1)*S bytes in per
//              GPU; best for latency-bound sizes.
//   two-shot : rank r reduces slice r only (reduce-scatter), runs Adadelta on that slice with its own
//              shard of the optimizer state, and stores the updated PARAMETERS into every peer
//              (all-gather of parameters, not of gradients).  2*(W-1)/W*S bytes per GPU.
//   NVLS     : two-shot whose reduce is one multimem.ld_reduce and whose parameter broadcast is one
//              multimem.st through the NVSwitch (multicast mapping from torch symmetric memory).
//
// Kernels of this file:
//   fused_allreduce_kernel<MODE, OPT>  the fc bucket (4.7 MB): entry barrier, reduce, Adadelta, parameters out.  Runs
//                                      on the engine's fc branch under the conv backward.
//   conv2_reduce_kernel<MODE>          conv2's part of the conv bucket: sums the wgrad kernel's partial rows into the
//                                      gradient, then chunk c is all-reduced by its owner (c % W) into EVERY rank's
//                                      "reduced" region R, followed by a per-chunk "R complete" flag in every rank.
//                                      Runs on the conv2-wgrad branch.
//   conv_finalize_kernel               end of the step: conv2 chunks wait for the LOCAL flag of their owner and apply
//                                      Adadelta from R; conv1's 5 chunks do a one-shot all-reduce behind the step's only
//                                      all-to-all barrier.
// No kernel ends with a barrier.  What makes that safe: a rank's gradient buckets are rewritten
// only by the next step's backward, which runs after that rank's finalize kernel, and the finalize kernel's all-to-all
// barrier cannot be passed before every rank has entered it, i.e. before every rank's fc-bucket and conv2-reduce
// kernels (earlier in that rank's graph) have finished reading peers.
//
// Barriers are flag words in the symmetric arena: CTA b of rank r writes a monotonically increasing
// token (derived from the device-side step counter, so the kernel is CUDA-graph replayable) into slot
// [b][r] of every peer with st.release.sys and spins with ld.acquire.sys on its own slots.  A wait is
// bounded in TIME (B200_COMM_TIMEOUT_MS, %globaltimer): on expiry the CTA fills in a CommFault record in
// mapped host memory (rank, peer, channel, expected and seen token) and traps, so the host can name the
// peer that went missing.  The kernels also run all ranks inside ONE cooperative launch
// (blockIdx.y = rank) so that they can be tested on a single GPU.
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "engine.h"
#include "symm.h"

namespace {

struct CommArgs {
  const float* grad[SYMM_MAX_WORLD];  // start of the bucket in every rank's gradient buffer
  float* param[SYMM_MAX_WORLD];       // start of the bucket in every rank's parameter (or output) buffer
  float* red[SYMM_MAX_WORLD];         // start of the bucket in every rank's reduced-gradient buffer (or null)
  uint32_t* flags[SYMM_MAX_WORLD];    // every rank's flag block
  const float* mc_grad;               // NVLS multicast aliases of the same three buffers (or null)
  float* mc_param;
  float* mc_red;
  CommFault* fault;                   // mapped pinned host record a timed-out barrier fills in before it traps (or null)
  unsigned long long timeout_ns;      // how long a barrier waits for a peer before giving up
  float* v;                           // local Adadelta state (bucket start); unused when OPT == false
  float* u;
  const B200Ctrl* ctrl;               // lr and step counter; may be null when token != 0 and OPT == false
  int rank, world, n4, channel;
  float inv_world;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {  // never served from a stale L1 line
  float4 r;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
// NVLS: one load returns the SUM over all ranks' copies (reduced inside the NVSwitch); one store is
// delivered to every rank's copy.
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Waits until *p has reached tok.  A peer that never arrives must not hang the GPU: after timeout_ns the
// waiting thread records who waited for whom in mapped host memory (the host prints it, symm.cu:
// b200_comm_fault_message) and traps, which ends the process with a launch failure instead of a hang.
__device__ __noinline__ void comm_wait_timeout(const CommArgs& a, int kind, int peer, uint32_t tok, uint32_t seen) {
  CommFault* f = a.fault;
  if (f && atomicCAS(&f->valid, 0u, 1u) == 0u) {
    f->rank = a.rank; f->world = a.world; f->channel = a.channel; f->kind = kind; f->block = blockIdx.x;
    f->peer = peer; f->expected = tok; f->seen = seen; f->step = a.ctrl ? a.ctrl->step : 0ull;
    f->waited_ns = a.timeout_ns;
    __threadfence_system();
    f->valid = 2u;
    __threadfence_system();
  }
  __trap();
}
__device__ __forceinline__ void comm_wait(const CommArgs& a, const uint32_t* p, uint32_t tok, int kind, int peer) {
  uint32_t seen = ld_acquire_sys(p);
  if ((int32_t)(seen - tok) >= 0) return;
  const unsigned long long t0 = global_ns();
  uint32_t spins = 0;
  while ((int32_t)((seen = ld_acquire_sys(p)) - tok) < 0) {
    if ((++spins & 1023u) == 0 && global_ns() - t0 > a.timeout_ns) comm_wait_timeout(a, kind, peer, tok, seen);
  }
}

// phase 0 = "my gradients are complete", phase 1 = "I have finished reading / my writes are out".
// No explicit fence is needed before the signal: gradients written by EARLIER kernels are visible
// system-wide at the kernel boundary, and the st.release.sys that follows the CTA barrier is cumulative
// over everything the CTA's threads wrote before that barrier (gradient chunks reduced in the same kernel,
// two-shot parameter stores).
__device__ __forceinline__ void comm_barrier(const CommArgs& a, int channel, int phase, uint32_t tok) {
  __syncthreads();
  if ((int)threadIdx.x < a.world) {
    const int idx = ((channel * 2 + phase) * SYMM_MAX_BLOCKS + blockIdx.x) * SYMM_MAX_WORLD;
    st_release_sys(a.flags[threadIdx.x] + idx + a.rank, tok);
    comm_wait(a, a.flags[a.rank] + idx + threadIdx.x, tok, phase, threadIdx.x);
  }
  __syncthreads();
}
// Gather form for chunk-owner kernels: every rank tells the OWNER of this CTA's chunk that its share is
// ready; only the owner waits.  (W-1) flag stores per chunk instead of W*(W-1).
__device__ __forceinline__ void comm_signal_owner(const CommArgs& a, int channel, int owner, uint32_t tok) {
  __syncthreads();
  const int idx = ((channel * 2) * SYMM_MAX_BLOCKS + blockIdx.x) * SYMM_MAX_WORLD;
  if (a.rank != owner) {
    if (threadIdx.x == 0) st_release_sys(a.flags[owner] + idx + a.rank, tok);
  } else {
    if ((int)threadIdx.x < a.world && (int)threadIdx.x != owner)
      comm_wait(a, a.flags[owner] + idx + threadIdx.x, tok, 2, threadIdx.x);
    __syncthreads();
  }
}

// advance > -1: the last CTA to finish also advances the control block (end-of-step marker)
// <= 64 registers/thread on purpose: this kernel runs on a side stream NEXT TO the persistent conv
// backward kernels and must not crowd their CTAs out of the register file.
template <int MODE, bool OPT>
__global__ void __launch_bounds__(512, 2)
fused_allreduce_kernel(const CommArgs* __restrict__ table, uint32_t token, float rho, float eps, int advance,
                       unsigned int* __restrict__ done_counter, int exit_barrier) {
  const CommArgs& a = table[blockIdx.y];
  const int W = a.world, r = a.rank;
  const uint32_t tok = token ? token : (uint32_t)(a.ctrl->step) * 2u + 1u;
  const float lr = OPT ? a.ctrl->lr : 0.f;
  const float inv_w = a.inv_world;
  const float* gp[SYMM_MAX_WORLD];
  float* pp[SYMM_MAX_WORLD];
#pragma unroll
  for (int q = 0; q < SYMM_MAX_WORLD; ++q) { gp[q] = a.grad[q < W ? q : 0]; pp[q] = a.param[q < W ? q : 0]; }

  int lo = 0, hi = a.n4;
  if (MODE >= 1) {
    const int per = (a.n4 + W - 1) / W;
    lo = min(a.n4, r * per);
    hi = min(a.n4, lo + per);
  }
  // Two elements per thread and round, all their peer loads in flight together: a peer load is a ~2 us round trip
  // over NVLink, and one element per round made the kernel a chain of such round trips (4 of them at W = 2).  The
  // local operands of the first round (parameters and Adadelta state of this rank's own slice, which no peer
  // touches before the barrier) are fetched BEFORE the barrier is waited for.
  constexpr int U = 2;
  const int stride = gridDim.x * blockDim.x;
  const int i0 = lo + blockIdx.x * blockDim.x + threadIdx.x;
  float4 p[U], vv[U], uu[U];
  auto load_state = [&](int i) {
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int e = i + k * stride;
      if (OPT && e < hi) {
        p[k] = *reinterpret_cast<const float4*>(a.param[r] + 4 * (size_t)e);
        vv[k] = *reinterpret_cast<const float4*>(a.v + 4 * (size_t)e);
        uu[k] = *reinterpret_cast<const float4*>(a.u + 4 * (size_t)e);
      }
    }
  };
  load_state(i0);

  comm_barrier(a, a.channel, 0, tok);

  for (int i = i0; i < hi; i += U * stride) {
    if (i != i0) load_state(i);
    float4 s[U];
#pragma unroll
    for (int k = 0; k < U; ++k) s[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (MODE == 2) {
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (i + k * stride < hi) s[k] = multimem_ld_reduce_f4(a.mc_grad + 4 * (size_t)(i + k * stride));  // in-switch reduction (NVLS)
    } else {
#pragma unroll
      for (int q = 0; q < SYMM_MAX_WORLD; ++q) {
        if (q < W) {  // fixed rank order on every rank: bit-identical sums
#pragma unroll
          for (int k = 0; k < U; ++k) {
            if (i + k * stride < hi) {
              const float4 g = ld_peer_f4(gp[q] + 4 * (size_t)(i + k * stride));
              s[k].x += g.x; s[k].y += g.y; s[k].z += g.z; s[k].w += g.w;
            }
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int e = i + k * stride;
      if (e >= hi) continue;
      float4 g = s[k];
      g.x *= inv_w; g.y *= inv_w; g.z *= inv_w; g.w *= inv_w;
      float4 out = g;
      if (OPT) {
        adadelta_update(g.x, p[k].x, vv[k].x, uu[k].x, lr, rho, eps);
        adadelta_update(g.y, p[k].y, vv[k].y, uu[k].y, lr, rho, eps);
        adadelta_update(g.z, p[k].z, vv[k].z, uu[k].z, lr, rho, eps);
        adadelta_update(g.w, p[k].w, vv[k].w, uu[k].w, lr, rho, eps);
        *reinterpret_cast<float4*>(a.v + 4 * (size_t)e) = vv[k];
        *reinterpret_cast<float4*>(a.u + 4 * (size_t)e) = uu[k];
        out = p[k];
      }
      if (MODE == 0) {
        *reinterpret_cast<float4*>(a.param[r] + 4 * (size_t)e) = out;
      } else if (MODE == 2) {
        multimem_st_f4(a.mc_param + 4 * (size_t)e, out);  // one store, delivered to every rank
      } else {
#pragma unroll
        for (int q = 0; q < SYMM_MAX_WORLD; ++q)
          if (q < W) st_peer_f4(pp[q] + 4 * (size_t)e, out);
      }
    }
  }
  if (MODE == 2) __threadfence_system();  // multimem stores must be performed before the exit signal
  // Inside the training step the exit barrier is redundant: the step's two communication kernels
  // alternate, and the ENTRY barrier of the next one (which every rank reaches only after this kernel
  // has completed on it) already proves that all peers are done reading / writing.  Stand-alone use
  // (microbenchmark, tests) keeps it.
  if (exit_barrier) comm_barrier(a, a.channel, 1, tok + 1u);
  if (advance >= 0 && threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(done_counter, 1u) == gridDim.x - 1) {  // every CTA has read ctrl->step / lr by now
      *done_counter = 0;
      B200Ctrl* c = const_cast<B200Ctrl*>(a.ctrl);
      c->step += 1;
      c->pos += advance;
    }
  }
}

// ---- the conv bucket (75 KB): two kernels ------------------------------------------------------------
// DDP's second bucket is {conv2.weight, conv2.bias, conv1.weight, conv1.bias}; its gradients become ready
// in two steps, and treating it as ONE all-reduce at the very end (as DDP does, and as round 1 of this
// code did) leaves the whole cross-GPU latency exposed after the last backward kernel.  Here:
//
//  conv2_reduce_kernel (on the conv2-wgrad branch, under conv2 dgrad / conv1 wgrad): one CTA per 64-float
//    chunk of conv2.{weight,bias}: sum the wgrad kernel's per-CTA partial rows into this rank's gradient
//    bucket, tell the chunk's owner, and the owner all-reduces the chunk into EVERY rank's reduced-gradient
//    buffer R (owner = chunk % W; P2P loads in fixed rank order + P2P stores, or multimem.ld_reduce +
//    multimem.st through the NVSwitch).  One-shot mode: every rank sums the chunk itself into its own R.
//    The parameters are not touched here: conv2 dgrad is still reading conv2.weight on every rank.
//    Each owner finally raises a per-chunk "R complete" flag on every rank.
//  conv_finalize_kernel (last kernel of the step): conv1's chunks (5 of 64 floats) are reduced from their
//    partial rows and exchanged one-shot behind an all-to-all barrier (every rank reads all W copies in rank
//    order: bit-identical sums, no second barrier); conv2's chunks only poll their LOCAL "R complete" flag and
//    read R; Adadelta runs on the whole conv bucket, and the last CTA advances the step counter.
// Ordering without exit barriers: the conv1 CTAs' all-to-all barrier is passed only when every rank has reached
// its finalize kernel, i.e. has finished its conv2_reduce and fc-bucket kernels, so all two-shot stores have landed;
// a rank reaches the NEXT step's communication kernels only after its own finalize kernel, so nobody
// overwrites a gradient chunk or R while a peer may still read the previous step's value.
// "chunk c of R is complete on this rank": written by the chunk's owner (or by the rank itself in one-shot mode) at
// the end of conv2_reduce_kernel, waited for by conv_finalize_kernel.  Phase-1 slot of the conv2 channel.
__device__ __forceinline__ int r_done_index(int channel, int chunk, int owner) {
  return ((channel * 2 + 1) * SYMM_MAX_BLOCKS + chunk) * SYMM_MAX_WORLD + owner;
}

template <int MODE>
__global__ void __launch_bounds__(FIN_THREADS)
conv2_reduce_kernel(const CommArgs* __restrict__ table, FinalizeSegs segs, const FinalizeSegs* __restrict__ segs_emul,
                    int channel) {
  __shared__ float4 red[FIN_SLICES][FIN_COLS];
  pdl_trigger();
  pdl_wait();
  const CommArgs& a = table[blockIdx.y];
  if (segs_emul) segs = segs_emul[blockIdx.y];
  const int W = a.world, r = a.rank;
  const uint32_t tok = (uint32_t)(a.ctrl->step) * 2u + 1u;
  const int col = threadIdx.x % FIN_COLS, off = blockIdx.x * FIN_CHUNK;
  const size_t i4 = (size_t)off + 4 * col;
  const float4 t = reduce_partial_chunk(segs, off, red);
  if (threadIdx.x < FIN_COLS) *reinterpret_cast<float4*>(const_cast<float*>(a.grad[r]) + i4) = t;
  if (MODE == 0) {
    comm_barrier(a, channel, 0, tok);
    if (threadIdx.x < FIN_COLS) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < SYMM_MAX_WORLD; ++q) {
        if (q < W) {  // fixed rank order on every rank: bit-identical sums
          const float4 g = ld_peer_f4(a.grad[q] + i4);
          s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
        }
      }
      s.x *= a.inv_world; s.y *= a.inv_world; s.z *= a.inv_world; s.w *= a.inv_world;
      *reinterpret_cast<float4*>(a.red[r] + i4) = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(a.flags[r] + r_done_index(channel, blockIdx.x, r), tok);  // R chunk complete here
  } else {
    const int owner = blockIdx.x % W;
    // The peers' chunks reach the owner's multimem load through their st.release.sys; the owner's own chunk was stored
    // just above through its unicast address, and same-address ordering does not extend to the multicast alias the
    // switch reads through: make it visible system-wide first.
    if (MODE == 2 && r == owner && threadIdx.x < FIN_COLS) __threadfence_system();
    comm_signal_owner(a, channel, owner, tok);
    if (r == owner && threadIdx.x < FIN_COLS) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      if (MODE == 2) {
        s = multimem_ld_reduce_f4(a.mc_grad + i4);  // in-switch reduction (NVLS)
      } else {
#pragma unroll
        for (int q = 0; q < SYMM_MAX_WORLD; ++q) {
          if (q < W) {
            const float4 g = ld_peer_f4(a.grad[q] + i4);
            s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
          }
        }
      }
      s.x *= a.inv_world; s.y *= a.inv_world; s.z *= a.inv_world; s.w *= a.inv_world;
      if (MODE == 2) {
        multimem_st_f4(a.mc_red + i4, s);  // one store, delivered to every rank
        __threadfence_system();
      } else {
#pragma unroll
        for (int q = 0; q < SYMM_MAX_WORLD; ++q)
          if (q < W) st_peer_f4(a.red[q] + i4, s);
      }
    }
    if (r == owner) {  // tell every rank that its copy of this chunk of R is complete (release: after the stores above)
      __syncthreads();
      if ((int)threadIdx.x < W) st_release_sys(a.flags[threadIdx.x] + r_done_index(channel, blockIdx.x, owner), tok);
    }
  }
}

// Last kernel of the step.  CTAs [0, n_c2) each take blockDim/FIN_COLS conv2 chunks: wait for the chunk's "R
// complete" flag (a LOCAL poll, normally long satisfied: conv2_reduce ran under the conv backward), read the reduced
// gradient from R, Adadelta.  The remaining CTAs own one conv1 chunk each: reduce its partial rows, exchange one-shot
// behind an all-to-all barrier, Adadelta.  Only those few CTAs talk to the peers at the end of the step (5 x (W-1)
// flag stores per rank instead of one barrier per chunk of the whole bucket).
__global__ void __launch_bounds__(FIN_THREADS)
conv_finalize_kernel(const CommArgs* __restrict__ table, FinalizeSegs segs, const FinalizeSegs* __restrict__ segs_emul,
                     int c2_chunks, int n_c2, int channel_c2, int owner_mod, float rho, float eps, int advance,
                     unsigned int* __restrict__ done_counter) {
  __shared__ float4 red[FIN_SLICES][FIN_COLS];
  pdl_trigger();
  pdl_wait();
  const CommArgs& a = table[blockIdx.y];
  if (segs_emul) segs = segs_emul[blockIdx.y];
  const int W = a.world, r = a.rank;
  const uint32_t tok = (uint32_t)(a.ctrl->step) * 2u + 1u;
  const float lr = a.ctrl->lr;
  const int col = threadIdx.x % FIN_COLS;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  size_t i4 = 0;
  bool active = false;
  if ((int)blockIdx.x < n_c2) {
    const int per = blockDim.x / FIN_COLS;
    const int chunk = blockIdx.x * per + threadIdx.x / FIN_COLS;
    active = chunk < c2_chunks;
    i4 = (size_t)chunk * FIN_CHUNK + 4 * col;
    if (active && col == 0) {
      const int owner = owner_mod ? chunk % W : r;
      comm_wait(a, a.flags[r] + r_done_index(channel_c2, chunk, owner), tok, 1, owner);
    }
    __syncwarp();  // the FIN_COLS = 16 lanes of a chunk sit in one warp; lane `col == 0` has acquired it for them
    if (active) s = ld_peer_f4(a.red[r] + i4);
  } else {
    const int chunk = c2_chunks + ((int)blockIdx.x - n_c2), off = chunk * FIN_CHUNK;
    i4 = (size_t)off + 4 * col;
    const float4 t = reduce_partial_chunk(segs, off, red);
    if (threadIdx.x < FIN_COLS) *reinterpret_cast<float4*>(const_cast<float*>(a.grad[r]) + i4) = t;
    comm_barrier(a, a.channel, 0, tok);  // release is cumulative over the gradient stores above
    active = threadIdx.x < FIN_COLS;
    if (active) {
#pragma unroll
      for (int q = 0; q < SYMM_MAX_WORLD; ++q) {
        if (q < W) {
          const float4 g = ld_peer_f4(a.grad[q] + i4);
          s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
        }
      }
      s.x *= a.inv_world; s.y *= a.inv_world; s.z *= a.inv_world; s.w *= a.inv_world;
    }
  }
  if (active) {
    float4 p = *reinterpret_cast<const float4*>(a.param[r] + i4);
    float4 vv = *reinterpret_cast<const float4*>(a.v + i4);
    float4 uu = *reinterpret_cast<const float4*>(a.u + i4);
    adadelta_update(s.x, p.x, vv.x, uu.x, lr, rho, eps);
    adadelta_update(s.y, p.y, vv.y, uu.y, lr, rho, eps);
    adadelta_update(s.z, p.z, vv.z, uu.z, lr, rho, eps);
    adadelta_update(s.w, p.w, vv.w, uu.w, lr, rho, eps);
    *reinterpret_cast<float4*>(a.v + i4) = vv;
    *reinterpret_cast<float4*>(a.u + i4) = uu;
    *reinterpret_cast<float4*>(a.param[r] + i4) = p;
  }
  if (advance >= 0) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(done_counter + blockIdx.y, 1u) == gridDim.x - 1) {  // every CTA has read ctrl->step / lr by now
        done_counter[blockIdx.y] = 0;
        B200Ctrl* c = const_cast<B200Ctrl*>(a.ctrl);
        c->step += 1;
        c->pos += advance;
      }
    }
  }
}

struct CommPlan {
  CommArgs* dev = nullptr;  // world_slots entries (1, or W when emulated)
  int slots = 1;
  int n4 = 0, world = 1;
  bool has_mc = false, has_red = false;
};

unsigned long long comm_timeout_ns() {
  // how long a barrier waits for a peer (B200_COMM_TIMEOUT_MS, default 60 s): long enough for rank-0-only work
  // between steps (evaluation, checkpoint writing, a blocked stdout pipe), short enough to end a dead job
  static const unsigned long long ns = [] {
    const char* e = getenv("B200_COMM_TIMEOUT_MS");
    const double ms = e ? atof(e) : 60000.0;
    return (unsigned long long)((ms > 0 ? ms : 60000.0) * 1e6);
  }();
  return ns;
}

// r_off: float offset of the reduced-gradient region inside the arena payload, or (size_t)-1 for none
int plan_from_arenas(SymmArena* const* arenas, int slots, size_t grad_off, size_t param_off, size_t v_off,
                     size_t u_off, size_t r_off, size_t n_floats, int channel, const B200Ctrl* const* ctrls,
                     CommPlan** out) {
  B200_REQUIRE(n_floats % 4 == 0 && grad_off % 4 == 0 && param_off % 4 == 0, "comm plan: offsets/size must be x4 floats");
  B200_REQUIRE(channel >= 0 && channel < SYMM_CHANNELS, "comm plan: channel %d out of range", channel);
  const bool has_red = r_off != (size_t)-1;
  std::vector<CommArgs> host(slots);
  for (int sidx = 0; sidx < slots; ++sidx) {
    SymmArena* a = arenas[sidx];
    B200_REQUIRE((grad_off + n_floats) * 4 <= a->payload && (param_off + n_floats) * 4 <= a->payload &&
                 (!has_red || (r_off + n_floats) * 4 <= a->payload), "comm plan: bucket exceeds the arena payload");
    CommArgs& c = host[sidx];
    memset(&c, 0, sizeof(c));
    for (int q = 0; q < a->world; ++q) {
      B200_REQUIRE(a->peer[q], "comm plan: rank %d is not mapped (import the IPC handles first)", q);
      c.grad[q] = (const float*)a->peer[q] + grad_off;
      c.param[q] = (float*)a->peer[q] + param_off;
      c.red[q] = has_red ? (float*)a->peer[q] + r_off : nullptr;
      c.flags[q] = a->flags(q);
    }
    c.mc_grad = a->mc ? (const float*)a->mc + grad_off : nullptr;
    c.mc_param = a->mc ? (float*)a->mc + param_off : nullptr;
    c.mc_red = (a->mc && has_red) ? (float*)a->mc + r_off : nullptr;
    c.fault = symm_fault_record(a);
    c.timeout_ns = comm_timeout_ns();
    c.v = (float*)a->peer[a->rank] + v_off;
    c.u = (float*)a->peer[a->rank] + u_off;
    c.ctrl = ctrls ? ctrls[sidx] : nullptr;
    c.rank = a->rank;
    c.world = a->world;
    c.n4 = (int)(n_floats / 4);
    c.channel = channel;
    c.inv_world = 1.0f / (float)a->world;
  }
  CommPlan* p = new CommPlan();
  p->slots = slots;
  p->n4 = (int)(n_floats / 4);
  p->world = arenas[0]->world;
  p->has_mc = arenas[0]->mc != nullptr;
  p->has_red = has_red;
  B200_CHECK(cudaMalloc(&p->dev, sizeof(CommArgs) * slots));
  B200_CHECK(cudaMemcpy(p->dev, host.data(), sizeof(CommArgs) * slots, cudaMemcpyHostToDevice));
  *out = p;
  return B200_OK;
}

int default_blocks(const CommPlan* p, int mode) {
  const int work = mode >= 1 ? (p->n4 + p->world - 1) / p->world : p->n4;
  int nb = (work + 1023) / 1024;  // ~2 float4 per thread
  if (nb < 1) nb = 1;
  const int cap = p->slots > 1 ? max(1, (b200_sm_count() * 2) / p->slots) : 64;
  return min(min(nb, cap), SYMM_MAX_BLOCKS);
}

int run_plan(const CommPlan* p, int mode, int opt, uint32_t token, int nblocks, float rho, float eps, cudaStream_t s,
             int exit_barrier = 1, int threads = 512) {
  B200_REQUIRE(p && mode >= 0 && mode <= 2, "comm run: bad plan/mode");
  B200_REQUIRE(mode != 2 || p->has_mc, "comm run: NVLS mode needs an arena with a multicast mapping");
  if (nblocks <= 0) nblocks = default_blocks(p, mode);
  B200_REQUIRE(nblocks <= SYMM_MAX_BLOCKS, "comm run: at most %d CTAs", SYMM_MAX_BLOCKS);
  dim3 grid(nblocks, p->slots);
  const CommArgs* tab = p->dev;
  int advance = -1;
  unsigned int* done_counter = nullptr;
  void* args[] = {(void*)&tab, (void*)&token, (void*)&rho, (void*)&eps, (void*)&advance, (void*)&done_counter,
                  (void*)&exit_barrier};
  const void* fn = mode == 0 ? (opt ? (const void*)fused_allreduce_kernel<0, true> : (const void*)fused_allreduce_kernel<0, false>)
                 : mode == 1 ? (opt ? (const void*)fused_allreduce_kernel<1, true> : (const void*)fused_allreduce_kernel<1, false>)
                             : (opt ? (const void*)fused_allreduce_kernel<2, true> : (const void*)fused_allreduce_kernel<2, false>);
  b200_prefer_max_shared(fn);  // so the side-stream reducer can share SMs with the persistent conv CTAs
  if (p->slots > 1) {
    // emulated ranks wait on each other inside one launch: all CTAs must be co-resident
    B200_CHECK(cudaLaunchCooperativeKernel(fn, grid, dim3(threads), args, 0, s));
  } else {
    B200_CHECK(cudaLaunchKernel(fn, grid, dim3(threads), args, 0, s));
  }
  b200_count_launch();
  return B200_OK;
}

// conv bucket, both kernels.  which: 0 = conv2_reduce (chunks [0, c2_chunks)), 1 = conv_finalize (all chunks).
// Emulated plans (slots > 1) launch cooperatively: W x chunks CTAs must be co-resident, so they use 64-thread CTAs.
int run_conv_bucket(const CommPlan* p, int which, int mode, const FinalizeSegs& segs, const FinalizeSegs* segs_emul,
                    int c2_chunks, int chunks, int channel_c2, int advance, unsigned int* done_counter, cudaStream_t s) {
  B200_REQUIRE(p && p->has_red, "conv bucket: the plan has no reduced-gradient region");
  B200_REQUIRE(mode >= 0 && mode <= 2 && (mode != 2 || p->has_mc), "conv bucket: bad mode %d", mode);
  B200_REQUIRE(chunks * FIN_COLS == p->n4 && c2_chunks > 0 && c2_chunks < chunks && chunks <= SYMM_MAX_BLOCKS,
               "conv bucket: %d chunks (%d of conv2) do not match a plan of %d float4", chunks, c2_chunks, p->n4);
  B200_REQUIRE(channel_c2 >= 0 && channel_c2 < SYMM_CHANNELS, "conv bucket: bad channel");
  const CommArgs* tab = p->dev;
  // emulated ranks: 4 slices per chunk instead of FIN_SLICES (the kernels derive the slice count from blockDim)
  const int threads = p->slots > 1 ? 4 * FIN_COLS : FIN_THREADS;
  FinalizeSegs sg = segs;
  float rho = 0.9f, eps = 1e-6f;
  if (which == 0) {
    dim3 grid(c2_chunks, p->slots);
    if (p->slots == 1) {
      if (mode == 0) B200_CHECK(b200_launch(conv2_reduce_kernel<0>, grid, dim3(threads), 0, s, tab, sg, segs_emul, channel_c2));
      else if (mode == 1) B200_CHECK(b200_launch(conv2_reduce_kernel<1>, grid, dim3(threads), 0, s, tab, sg, segs_emul, channel_c2));
      else B200_CHECK(b200_launch(conv2_reduce_kernel<2>, grid, dim3(threads), 0, s, tab, sg, segs_emul, channel_c2));
    } else {
      void* args[] = {(void*)&tab, (void*)&sg, (void*)&segs_emul, (void*)&channel_c2};
      const void* fn = mode == 0 ? (const void*)conv2_reduce_kernel<0>
                     : mode == 1 ? (const void*)conv2_reduce_kernel<1> : (const void*)conv2_reduce_kernel<2>;
      B200_CHECK(cudaLaunchCooperativeKernel(fn, grid, dim3(threads), args, 0, s));
    }
  } else {
    const int per = threads / FIN_COLS;
    int n_c2 = (c2_chunks + per - 1) / per, owner_mod = mode == 0 ? 0 : 1;
    dim3 grid(n_c2 + (chunks - c2_chunks), p->slots);
    if (p->slots == 1) {
      B200_CHECK(b200_launch(conv_finalize_kernel, grid, dim3(threads), 0, s, tab, sg, segs_emul, c2_chunks, n_c2, channel_c2,
                             owner_mod, rho, eps, advance, done_counter));
    } else {
      void* args[] = {(void*)&tab, (void*)&sg, (void*)&segs_emul, (void*)&c2_chunks, (void*)&n_c2, (void*)&channel_c2,
                      (void*)&owner_mod, (void*)&rho, (void*)&eps, (void*)&advance, (void*)&done_counter};
      B200_CHECK(cudaLaunchCooperativeKernel((const void*)conv_finalize_kernel, grid, dim3(threads), args, 0, s));
    }
  }
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ---- engine hooks ---------------------------------------------------------------------------------
constexpr int CH_FC = 0, CH_CONV = 1, CH_CONV2 = 3;  // barrier channels (2 = stand-alone calls: tests, microbenchmark)
struct DdpState {
  CommPlan* plan[2] = {nullptr, nullptr};
  int mode[2] = {0, 0};
};

int hook_bucket_ready(Engine* E, int bucket, cudaStream_t s) {
  DdpState* d = (DdpState*)E->hooks.state;
  if (bucket == 0) {
    // The fc bucket is complete while the conv backward still runs; `s` is the engine's fc branch.  No exit
    // barrier: the finalize kernel's entry barrier orders this kernel's peer accesses (see above).
    // Footprint of the reducer (env-tunable).  Measured at W = 2 with the state prefetch in place: 74x512 ... 296x128
    // all within 148.6 - 150.9 us/step, i.e. no effect; one CTA per SM is the default.
    static const int nb0 = getenv("B200_DDP_B0_BLOCKS") ? atoi(getenv("B200_DDP_B0_BLOCKS")) : 148;
    static const int th0 = getenv("B200_DDP_B0_THREADS") ? atoi(getenv("B200_DDP_B0_THREADS")) : 256;
    static const int exit_b = getenv("B200_DDP_EXIT_BARRIER") ? atoi(getenv("B200_DDP_EXIT_BARRIER")) : 0;
    return run_plan(d->plan[0], d->mode[0], 1, 0, min(nb0, SYMM_MAX_BLOCKS), 0.9f, 1e-6f, s, exit_b, th0);
  }
  if (bucket == 2)  // conv2's partial rows are complete; `s` is the engine's conv2-wgrad branch
    return run_conv_bucket(d->plan[1], 0, d->mode[1], E->fin_c2, nullptr, E->c2_chunks, E->conv_chunks, CH_CONV2, -1,
                           nullptr, s);
  return B200_OK;
}

int hook_step_end(Engine* E, int advance, cudaStream_t s) {
  DdpState* d = (DdpState*)E->hooks.state;
  return run_conv_bucket(d->plan[1], 1, d->mode[1], E->fin_c1, nullptr, E->c2_chunks, E->conv_chunks, CH_CONV2, advance,
                         E->done_counter, s);
}

void hook_destroy(Engine* E) {
  DdpState* d = (DdpState*)E->hooks.state;
  if (!d) return;
  for (int k = 0; k < 2; ++k)
    if (d->plan[k]) { cudaFree(d->plan[k]->dev); delete d->plan[k]; }
  delete d;
  E->hooks.state = nullptr;
}

}  // namespace

// ---- C ABI --------------------------------------------------------------------------------------
// A plan binds one bucket (offsets in floats inside the arena payload) to a barrier channel.
extern "C" int b200_comm_plan_create(void* arena, size_t grad_off, size_t param_off, size_t v_off, size_t u_off,
                                     size_t n_floats, int channel, const void* ctrl, void** plan_out) {
  B200_REQUIRE(arena && plan_out, "comm_plan_create: null argument");
  SymmArena* a = (SymmArena*)arena;
  const B200Ctrl* c = (const B200Ctrl*)ctrl;
  CommPlan* p = nullptr;
  int r = plan_from_arenas(&a, 1, grad_off, param_off, v_off, u_off, (size_t)-1, n_floats, channel, ctrl ? &c : nullptr, &p);
  if (r != B200_OK) return r;
  *plan_out = p;
  return B200_OK;
}

// all `world` emulated arenas at once; ctrls: world control blocks (or NULL); r_off: reduced-gradient region or -1
extern "C" int b200_comm_plan_create_emulated(void** arenas, int world, size_t grad_off, size_t param_off, size_t v_off,
                                              size_t u_off, size_t n_floats, int channel, const void** ctrls,
                                              void** plan_out) {
  B200_REQUIRE(arenas && plan_out && world >= 1 && world <= SYMM_MAX_WORLD, "comm_plan_create_emulated: bad arguments");
  CommPlan* p = nullptr;
  int r = plan_from_arenas((SymmArena* const*)arenas, world, grad_off, param_off, v_off, u_off, (size_t)-1, n_floats,
                           channel, (const B200Ctrl* const*)ctrls, &p);
  if (r != B200_OK) return r;
  *plan_out = p;
  return B200_OK;
}

// Emulated conv-bucket plan for single-GPU tests of conv2_reduce_kernel / conv_finalize_kernel: payload layout of
// every arena in floats [params n][grads n][sq_avg n][acc_delta n][reduced n], n = 256 * chunks.
extern "C" int b200_comm_conv_plan_create_emulated(void** arenas, int world, size_t n_floats, const void** ctrls,
                                                   void** plan_out) {
  B200_REQUIRE(arenas && plan_out && ctrls && world >= 1 && world <= SYMM_MAX_WORLD, "comm_conv_plan_create_emulated: bad arguments");
  CommPlan* p = nullptr;
  int r = plan_from_arenas((SymmArena* const*)arenas, world, n_floats, 0, 2 * n_floats, 3 * n_floats, 4 * n_floats, n_floats,
                           CH_CONV, (const B200Ctrl* const*)ctrls, &p);
  if (r != B200_OK) return r;
  *plan_out = p;
  return B200_OK;
}

// One emulated conv-bucket exchange: conv2_reduce (mode 0 one-shot / 1 two-shot) then conv_finalize.
//   c2_parts[r]: rank r's conv2 partial rows [np2][18496]; c1_parts[r]: its conv1 group rows [ng][320];
//   counters: >= world zero-initialised words.  The conv bucket layout is the engine's:
//   [conv2.weight 18432 | conv2.bias 256 | conv1.weight 512 | conv1.bias 256].
extern "C" int b200_comm_conv_bucket_emulated(void* plan, int mode, const float* const* c2_parts, int np2,
                                              const float* const* c1_parts, int ng, int advance, unsigned int* counters,
                                              void* stream) {
  CommPlan* p = (CommPlan*)plan;
  B200_REQUIRE(p && c2_parts && c1_parts && counters && p->slots >= 1, "comm_conv_bucket_emulated: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<FinalizeSegs> h2(p->slots), h1(p->slots);
  for (int r = 0; r < p->slots; ++r) {
    h2[r].seg[0] = {c2_parts[r], np2, 18496, 18432, 0};
    h2[r].seg[1] = {c2_parts[r] + 18432, np2, 18496, 64, 18432};
    h1[r].seg[0] = {c1_parts[r], ng, 320, 288, 18688};
    h1[r].seg[1] = {c1_parts[r] + 288, ng, 320, 32, 19200};
  }
  FinalizeSegs* dev = nullptr;
  B200_CHECK(cudaMalloc(&dev, sizeof(FinalizeSegs) * 2 * p->slots));
  B200_CHECK(cudaMemcpyAsync(dev, h2.data(), sizeof(FinalizeSegs) * p->slots, cudaMemcpyHostToDevice, s));
  B200_CHECK(cudaMemcpyAsync(dev + p->slots, h1.data(), sizeof(FinalizeSegs) * p->slots, cudaMemcpyHostToDevice, s));
  int r = run_conv_bucket(p, 0, mode, h2[0], dev, 18688 / FIN_CHUNK, 19456 / FIN_CHUNK, CH_CONV2, -1, nullptr, s);
  if (r == B200_OK) r = run_conv_bucket(p, 1, mode, h1[0], dev + p->slots, 18688 / FIN_CHUNK, 19456 / FIN_CHUNK, CH_CONV2, advance, counters, s);
  cudaStreamSynchronize(s);
  cudaFree(dev);
  return r;
}

extern "C" int b200_comm_plan_destroy(void* plan) {
  CommPlan* p = (CommPlan*)plan;
  if (!p) return B200_OK;
  cudaFree(p->dev);
  delete p;
  return B200_OK;
}

// mode 0 one-shot, 1 two-shot, 2 NVLS.  opt 1: fused 1/W + Adadelta (needs ctrl in the plan), 0: plain averaged
// all-reduce into the `param_off` buffer.  token 0: derive the barrier token from ctrl->step; otherwise
// the caller supplies a token that grows by 2 from call to call.  nblocks 0: pick automatically.
extern "C" int b200_comm_run(void* plan, int mode, int opt, unsigned int token, int nblocks, void* stream) {
  return run_plan((const CommPlan*)plan, mode, opt, token, nblocks, 0.9f, 1e-6f, (cudaStream_t)stream);
}

// Installs the data-parallel hooks on an engine whose params/grads/state live in `arena`
// (payload layout in floats: [params T][grads T][sq_avg T][acc_delta T][reduced conv gradients T]).
extern "C" int b200_ddp_attach(void* engine, void* arena, int mode_bucket0, int mode_bucket1) {
  Engine* E = (Engine*)engine;
  SymmArena* a = (SymmArena*)arena;
  B200_REQUIRE(E && a, "ddp_attach: null argument");
  const size_t T = (size_t)E->cfg.param_floats;
  B200_REQUIRE(a->payload >= 5 * T * 4, "ddp_attach: arena too small for params+grads+state+reduced gradients");
  B200_REQUIRE((float*)a->peer[a->rank] == E->P && E->G[0] == E->P + T && E->V == E->P + 2 * T && E->U == E->P + 3 * T,
               "ddp_attach: the engine must have been created on this arena's buffers");
  B200_REQUIRE(mode_bucket0 >= 0 && mode_bucket0 <= 2 && mode_bucket1 >= 0 && mode_bucket1 <= 2, "ddp_attach: bad mode");
  B200_REQUIRE((mode_bucket0 != 2 && mode_bucket1 != 2) || a->mc, "ddp_attach: NVLS mode needs a multicast mapping");
  DdpState* d = new DdpState();
  d->mode[0] = mode_bucket0;
  d->mode[1] = mode_bucket1;
  const B200Ctrl* ctrl = E->ctrl;
  for (int k = 0; k < 2; ++k) {
    const size_t beg = (size_t)E->cfg.bucket_begin[k], n = (size_t)(E->cfg.bucket_end[k] - E->cfg.bucket_begin[k]);
    int r = plan_from_arenas(&a, 1, T + beg, beg, 2 * T + beg, 3 * T + beg, k == 1 ? 4 * T + beg : (size_t)-1, n,
                             k == 0 ? CH_FC : CH_CONV, &ctrl, &d->plan[k]);
    if (r != B200_OK) {
      for (int j = 0; j < k; ++j) { cudaFree(d->plan[j]->dev); delete d->plan[j]; }
      delete d;
      return r;
    }
  }
  E->hooks.state = d;
  E->hooks.bucket_ready = hook_bucket_ready;
  E->hooks.step_end = hook_step_end;
  E->hooks.destroy = hook_destroy;
  return B200_OK;
}

// copy `n_floats` starting at float offset `off` from rank `src`'s arena into the local one (init only)
__global__ void peer_copy_kernel(float4* __restrict__ dst, const float4* __restrict__ src, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}
extern "C" int b200_symm_pull(void* arena, int src, size_t off, size_t n_floats, void* stream) {
  SymmArena* a = (SymmArena*)arena;
  B200_REQUIRE(a && src >= 0 && src < a->world && off % 4 == 0 && n_floats % 4 == 0, "symm_pull: bad arguments");
  if (src == a->rank) return B200_OK;
  peer_copy_kernel<<<64, 512, 0, (cudaStream_t)stream>>>((float4*)((float*)a->peer[a->rank] + off),
                                                        (const float4*)((const float*)a->peer[src] + off), n_floats / 4);
  b200_count_launch();
  B200_LAUNCH_CHECK();
  return B200_OK;
}
