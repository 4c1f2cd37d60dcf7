// Symmetric memory: one device allocation per rank, mapped into every peer with CUDA IPC, so that
// kernels on any GPU can load/store every rank's gradient and parameter buckets directly over
// NVLink / NVSwitch.  This is the substrate of the fused all-reduce + Adadelta kernels
// (allreduce.cu) that replace DistributedDataParallel's bucket copies and ncclAllReduce calls
// (mnist_ddp.py:173).
//
// Bootstrap stays with torch.distributed: each rank creates its arena, exports a 64-byte IPC handle,
// the handles are all-gathered through the process group, and every rank imports its peers'.
// Layout of an arena: [payload bytes][flag block]; the flag block holds the per-bucket, per-CTA,
// per-source-rank barrier words that the kernels signal with st.release.sys / ld.acquire.sys.
//
// For tests on a single GPU an "emulated" set of W arenas can be created inside one process: the
// same kernels then run all W ranks in one cooperative launch (blockIdx.y = rank).
#include <cstring>

#include "common.cuh"
#include "symm.h"

CommFault* symm_fault_record(SymmArena* a) {
  if (a->fault_dev) return a->fault_dev;
  void* h = nullptr;
  if (cudaHostAlloc(&h, sizeof(CommFault), cudaHostAllocMapped) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  memset(h, 0, sizeof(CommFault));
  void* d = nullptr;
  if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) { cudaGetLastError(); cudaFreeHost(h); return nullptr; }
  a->fault_host = (CommFault*)h;
  a->fault_dev = (CommFault*)d;
  b200_register_fault_record(a->fault_host);
  return a->fault_dev;
}

extern "C" int b200_symm_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
extern "C" int b200_symm_flag_bytes(void) { return SYMM_FLAG_BYTES; }

extern "C" int b200_symm_create(int world, int rank, size_t payload_bytes, void** out) {
  B200_REQUIRE(out && world >= 1 && world <= SYMM_MAX_WORLD && rank >= 0 && rank < world,
               "symm_create: bad world/rank %d/%d (max world %d)", world, rank, SYMM_MAX_WORLD);
  SymmArena* a = new SymmArena();
  a->world = world;
  a->rank = rank;
  a->payload = (payload_bytes + 255) & ~(size_t)255;
  B200_CHECK(cudaGetDevice(&a->device));
  B200_CHECK(cudaMalloc(&a->peer[rank], a->payload + SYMM_FLAG_BYTES));
  B200_CHECK(cudaMemset(a->peer[rank], 0, a->payload + SYMM_FLAG_BYTES));
  a->owns[rank] = true;
  B200_CHECK(cudaDeviceSynchronize());
  *out = a;
  return B200_OK;
}

extern "C" int b200_symm_export(void* arena, void* handle_out) {
  SymmArena* a = (SymmArena*)arena;
  B200_REQUIRE(a && handle_out, "symm_export: null argument");
  cudaIpcMemHandle_t h;
  B200_CHECK(cudaIpcGetMemHandle(&h, a->peer[a->rank]));
  memcpy(handle_out, &h, sizeof(h));
  return B200_OK;
}

// all_handles: world consecutive handles, indexed by rank (own slot ignored)
extern "C" int b200_symm_import(void* arena, const void* all_handles) {
  SymmArena* a = (SymmArena*)arena;
  B200_REQUIRE(a && all_handles, "symm_import: null argument");
  for (int q = 0; q < a->world; ++q) {
    if (q == a->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)all_handles + (size_t)q * sizeof(h), sizeof(h));
    B200_CHECK(cudaIpcOpenMemHandle(&a->peer[q], h, cudaIpcMemLazyEnablePeerAccess));
    a->opened[q] = true;
  }
  return B200_OK;
}

// Wrap memory that somebody else made symmetric (e.g. torch.distributed._symmetric_memory, which also
// provides an NVLS multicast mapping): peer_ptrs[q] = rank q's buffer as mapped here, mc_ptr = the
// multicast alias or 0.  The buffers must hold payload_bytes + b200_symm_flag_bytes() and the flag
// block must be zero.  Nothing is owned: b200_symm_destroy only frees the descriptor.
extern "C" int b200_symm_wrap(int world, int rank, size_t payload_bytes, const uint64_t* peer_ptrs, uint64_t mc_ptr,
                              void** out) {
  B200_REQUIRE(out && peer_ptrs && world >= 1 && world <= SYMM_MAX_WORLD && rank >= 0 && rank < world,
               "symm_wrap: bad arguments");
  B200_REQUIRE(payload_bytes % 256 == 0, "symm_wrap: payload must be a multiple of 256 bytes");
  SymmArena* a = new SymmArena();
  a->world = world;
  a->rank = rank;
  a->payload = payload_bytes;
  B200_CHECK(cudaGetDevice(&a->device));
  for (int q = 0; q < world; ++q) a->peer[q] = (void*)peer_ptrs[q];
  a->mc = (void*)mc_ptr;
  *out = a;
  return B200_OK;
}

extern "C" int b200_symm_has_multicast(void* arena) { return arena && ((SymmArena*)arena)->mc ? 1 : 0; }

// W arenas on the current device, cross-linked as if they were W ranks (single-process tests)
extern "C" int b200_symm_create_emulated(int world, size_t payload_bytes, void** arenas_out) {
  B200_REQUIRE(arenas_out && world >= 1 && world <= SYMM_MAX_WORLD, "symm_create_emulated: bad world %d", world);
  SymmArena* a[SYMM_MAX_WORLD];
  for (int r = 0; r < world; ++r) {
    void* p = nullptr;
    int rc = b200_symm_create(world, r, payload_bytes, &p);
    if (rc != B200_OK) return rc;
    a[r] = (SymmArena*)p;
    a[r]->emulated = true;
  }
  for (int r = 0; r < world; ++r)
    for (int q = 0; q < world; ++q) a[r]->peer[q] = a[q]->peer[q];
  for (int r = 0; r < world; ++r) arenas_out[r] = a[r];
  return B200_OK;
}

extern "C" int b200_symm_peer(void* arena, int q, void** ptr) {
  SymmArena* a = (SymmArena*)arena;
  B200_REQUIRE(a && ptr && q >= 0 && q < a->world && a->peer[q], "symm_peer: rank %d not mapped", q);
  *ptr = a->peer[q];
  return B200_OK;
}

extern "C" int b200_symm_destroy(void* arena) {
  SymmArena* a = (SymmArena*)arena;
  if (!a) return B200_OK;
  cudaDeviceSynchronize();
  for (int q = 0; q < a->world; ++q)
    if (a->opened[q]) cudaIpcCloseMemHandle(a->peer[q]);
  if (a->owns[a->rank]) cudaFree(a->peer[a->rank]);
  if (a->fault_host) { b200_unregister_fault_record(a->fault_host); cudaFreeHost(a->fault_host); }
  delete a;
  return B200_OK;
}
