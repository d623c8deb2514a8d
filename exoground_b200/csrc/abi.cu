// abi.cu — process-wide plumbing of the C ABI: version, thread-local error text, launch counter.
#include "common.cuh"
#include <atomic>
#include <string.h>

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void exg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void exg_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// SMs set aside for kernels of other streams (NCCL during an overlapped gradient all-reduce): persistent kernels
// size their grid to the remaining SMs, so none of their CTAs queues behind a communication CTA.
static std::atomic<int> g_reserved_sms{0};
int exg_persistent_sms() {
  const int n = exg_num_sms() - g_reserved_sms.load(std::memory_order_relaxed);
  return n < 2 ? 2 : n;
}
extern "C" int exg_reserve_sms(int n) {
  if (n < 0) n = 0;
  return g_reserved_sms.exchange(n, std::memory_order_relaxed);
}

extern "C" int exg_version(void) { return 100; }
extern "C" const char* exg_last_error(void) { return g_err; }
extern "C" long long exg_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

// ---------------------------------------------------------------- launch profiling
#include <mutex>
#include <vector>
namespace {
struct ProfRec { int cls; double work; cudaEvent_t e0, e1; };
std::mutex g_prof_mu;
bool g_prof_on = false;
std::vector<ProfRec> g_prof_recs;
std::vector<cudaEvent_t> g_prof_pool;
const char* kProfNames[EXG_PC_COUNT] = {"gemm_tc", "gemm_simt", "attn_fwd", "attn_bwd", "layernorm_fwd",
                                        "layernorm_bwd", "rowwise", "nce_fwd", "nce_bwd", "optim"};
cudaEvent_t prof_event() {
  if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
  cudaEvent_t e; cudaEventCreate(&e); return e;
}
}  // namespace
bool exg_profiling() { return g_prof_on; }
void exg_prof_open(int cls, double work, cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r; r.cls = cls; r.work = work; r.e0 = prof_event(); r.e1 = prof_event();
  cudaEventRecord(r.e0, s);
  g_prof_recs.push_back(r);
}
void exg_prof_close(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_prof_recs.empty()) cudaEventRecord(g_prof_recs.back().e1, s);
}
extern "C" int exg_profile_classes(void) { return EXG_PC_COUNT; }
extern "C" const char* exg_profile_class_name(int cls) {
  return (cls >= 0 && cls < EXG_PC_COUNT) ? kProfNames[cls] : "";
}
extern "C" int exg_profile_begin(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = true;
  return 0;
}
extern "C" int exg_profile_end(double* ms, double* work, long long* launches) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = false;
  for (int c = 0; c < EXG_PC_COUNT; ++c) { ms[c] = 0; work[c] = 0; launches[c] = 0; }
  int rc = 0;
  for (auto& r : g_prof_recs) {
    float t = 0.f;
    cudaError_t e = cudaEventSynchronize(r.e1);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&t, r.e0, r.e1);
    if (e != cudaSuccess) { exg_set_error("profile_end: %s", cudaGetErrorString(e)); rc = (int)e; }
    ms[r.cls] += t; work[r.cls] += r.work; launches[r.cls] += 1;
    g_prof_pool.push_back(r.e0); g_prof_pool.push_back(r.e1);
  }
  g_prof_recs.clear();
  return rc;
}
