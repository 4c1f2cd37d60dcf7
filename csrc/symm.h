// Symmetric-memory arena shared by symm.cu (set-up) and allreduce.cu (kernels, engine hooks).
#pragma once
#include <stddef.h>
#include <stdint.h>

constexpr int SYMM_MAX_WORLD = 8;
constexpr int SYMM_MAX_BLOCKS = 512;   // CTAs per communication kernel
constexpr int SYMM_CHANNELS = 4;       // independent barrier channels (fc bucket, conv bucket, stand-alone calls, conv2 half)
// flag word index: ((channel * 2 + phase) * SYMM_MAX_BLOCKS + block) * SYMM_MAX_WORLD + source_rank
constexpr int SYMM_FLAG_WORDS = SYMM_CHANNELS * 2 * SYMM_MAX_BLOCKS * SYMM_MAX_WORLD;
constexpr int SYMM_FLAG_BYTES = SYMM_FLAG_WORDS * 4;

struct SymmArena {
  int world = 1, rank = 0, device = 0;
  size_t payload = 0;                      // bytes before the flag block
  void* peer[SYMM_MAX_WORLD] = {nullptr};  // base address of every rank's arena as mapped in THIS process
  bool opened[SYMM_MAX_WORLD] = {false};
  bool owns[SYMM_MAX_WORLD] = {false};
  bool emulated = false;
  void* mc = nullptr;                      // multicast (NVLS) alias of all ranks' arenas, or null
  struct CommFault* fault_host = nullptr;  // mapped pinned record a timed-out barrier fills in (common.cuh)
  struct CommFault* fault_dev = nullptr;   // its device alias
  uint32_t* flags(int q) const { return (uint32_t*)((char*)peer[q] + payload); }
};

// device alias of the arena's barrier-fault record (allocated on first use; null if mapped memory is unavailable)
struct CommFault* symm_fault_record(SymmArena* a);
