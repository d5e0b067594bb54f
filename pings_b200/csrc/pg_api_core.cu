// Core of the C-ABI library: version, error string, and the opt-in per-phase device timer
// bench.py uses to attribute a step to kernels with CUDA events on the launching stream.
#include <vector>
#include "pg_common.cuh"

namespace pg {
thread_local char g_last_error[256] = {0};

struct PhaseRec {
  const char* name;
  cudaEvent_t a, b;
};
static thread_local bool g_timing = false;
static thread_local std::vector<PhaseRec> g_phases;
static thread_local std::vector<cudaEvent_t> g_event_pool;

static cudaEvent_t get_event() {
  if (!g_event_pool.empty()) {
    cudaEvent_t e = g_event_pool.back();
    g_event_pool.pop_back();
    return e;
  }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}

bool timing_on() { return g_timing; }
void phase_begin(const char* name, cudaStream_t st) {
  if (!g_timing) return;
  PhaseRec r{name, get_event(), get_event()};
  cudaEventRecord(r.a, st);
  g_phases.push_back(r);
}
void phase_end(cudaStream_t st) {
  if (!g_timing || g_phases.empty()) return;
  cudaEventRecord(g_phases.back().b, st);
}
}  // namespace pg

using namespace pg;

extern "C" {

int pg_abi_version(void) { return 3; }
const char* pg_last_error(void) { return g_last_error; }

int pg_timing_enable(int on) {
  g_timing = on != 0;
  return PG_OK;
}

// Drains the phases recorded on this thread since the last call: phase i ran for ms[i] and is named
// names[i] (static strings). Returns the number of phases written (<= max_phases), or -1 on error.
int pg_timing_drain(int max_phases, const char** names, float* ms) {
  int n = 0;
  for (auto& r : g_phases) {
    float t = 0.f;
    if (cudaEventSynchronize(r.b) != cudaSuccess || cudaEventElapsedTime(&t, r.a, r.b) != cudaSuccess) {
      g_phases.clear();
      return -1;
    }
    if (n < max_phases) {
      names[n] = r.name;
      ms[n] = t;
      ++n;
    }
    g_event_pool.push_back(r.a);
    g_event_pool.push_back(r.b);
  }
  g_phases.clear();
  return n;
}

}  // extern "C"
