// Process-wide runtime state of libb200mnist: error strings, launch accounting, device facts.
#include <atomic>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_set>
#include <vector>

#include "common.cuh"

static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

static std::mutex g_fault_mu;
static std::vector<CommFault*> g_fault_records;

void b200_register_fault_record(CommFault* p) {
  std::lock_guard<std::mutex> lock(g_fault_mu);
  g_fault_records.push_back(p);
}
void b200_unregister_fault_record(CommFault* p) {
  std::lock_guard<std::mutex> lock(g_fault_mu);
  for (size_t i = 0; i < g_fault_records.size(); ++i)
    if (g_fault_records[i] == p) { g_fault_records.erase(g_fault_records.begin() + i); break; }
}

// text of the first complete cross-GPU barrier fault record, or 0 characters
static int fault_text(char* buf, size_t len) {
  std::lock_guard<std::mutex> lock(g_fault_mu);
  for (CommFault* f : g_fault_records) {
    if (!f || *(volatile unsigned int*)&f->valid != 2u) continue;
    static const char* kinds[] = {"entry barrier", "exit barrier", "chunk owner waiting for contributions"};
    return snprintf(buf, len,
                    "cross-GPU barrier timed out: rank %d of %d waited %.1f s for rank %d (channel %d, %s, CTA %d, step "
                    "%llu, expected token %u, last seen %u) -- that peer exited, hung, or is more than "
                    "B200_COMM_TIMEOUT_MS behind", f->rank, f->world, (double)f->waited_ns * 1e-9, f->peer, f->channel,
                    kinds[f->kind >= 0 && f->kind < 3 ? f->kind : 0], f->block, f->step, f->expected, f->seen);
  }
  return 0;
}

void b200_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  int n = vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  if (n < 0) n = 0;
  if ((size_t)n >= sizeof(g_err) - 8) return;
  char extra[512];
  if (fault_text(extra, sizeof(extra)) > 0) snprintf(g_err + n, sizeof(g_err) - n, " [%s]", extra);
}

// the pending barrier-fault message (empty string when no barrier has timed out); for hosts that see a CUDA error
// from some other library first (e.g. torch) and want to know whether a lost peer is behind it
extern "C" const char* b200_comm_fault_message(void) {
  static thread_local char buf[512];
  buf[0] = 0;
  fault_text(buf, sizeof(buf));
  return buf;
}

void b200_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool b200_pdl_enabled(const void* kernel) {
  // B200_PDL=1 turns programmatic dependent launch on; B200_PDL_MASK (diagnostics) limits it to the kernels
  // whose first-launch index has its bit set.
  static const bool on = getenv("B200_PDL") && atoi(getenv("B200_PDL")) == 1;
  if (!on) return false;
  static const long mask = getenv("B200_PDL_MASK") ? strtol(getenv("B200_PDL_MASK"), nullptr, 0) : -1;
  if (mask == -1) return true;
  static std::mutex mu;
  static std::vector<const void*> order;
  std::lock_guard<std::mutex> lock(mu);
  size_t i = 0;
  for (; i < order.size(); ++i) if (order[i] == kernel) break;
  if (i == order.size()) order.push_back(kernel);
  return i < 63 && ((mask >> i) & 1);
}

void b200_prefer_max_shared(const void* kernel) {
  static std::mutex mu;
  static std::unordered_set<const void*> seen;
  std::lock_guard<std::mutex> lock(mu);
  if (!seen.insert(kernel).second) return;
  static const bool off = getenv("B200_NO_CARVEOUT") != nullptr;  // diagnostics
  // only a scheduling hint: a refusal must not surface as a launch error later
  if (!off && cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                   cudaSharedmemCarveoutMaxShared) != cudaSuccess)
    cudaGetLastError();
}

int b200_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = n > 0 ? n : 148;
  }
  return cached[dev];
}

__global__ void ctrl_set_kernel(B200Ctrl* c, unsigned long long step, unsigned long long seed, int pos, int train,
                                float lr, float inv_world) {
  c->step = step;
  c->seed = seed;
  c->pos = pos;
  c->train = train;
  c->lr = lr;
  c->inv_world = inv_world;
}

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_version(void) { return 100; }
extern "C" long long b200_launch_count(void) { return g_launches.load(); }
extern "C" int b200_ctrl_bytes(void) { return (int)sizeof(B200Ctrl); }

extern "C" int b200_ctrl_set(void* ctrl, unsigned long long step, unsigned long long seed, int pos, int train,
                             float lr, float inv_world, void* stream) {
  B200_REQUIRE(ctrl != nullptr, "ctrl_set: null ctrl");
  ctrl_set_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((B200Ctrl*)ctrl, step, seed, pos, train, lr, inv_world);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
