// Error plumbing and version for the C ABI.
#include "../../include/vistaf.h"
#include "common.cuh"
#include <string.h>

char g_vistaf_err[512] = "";

int vistaf_set_err(const char* what, cudaError_t e, const char* file, int line) {
  snprintf(g_vistaf_err, sizeof(g_vistaf_err), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  return 1;
}

extern "C" const char* vistaf_last_error(void) { return g_vistaf_err; }
extern "C" int vistaf_version(void) { return 100; }

// ------------------------------------------------------------------ stage timers (CUDA events on the launch stream)
// Off by default (no events recorded).  When enabled, every VISTAF_STAGE(name, stream) mark records an event;
// the time between two consecutive marks is charged to the earlier mark's name.  bench.py reads the totals to
// report which stage dominates and its achieved bandwidth, measured live on the stream the kernels run on.
#include <vector>
#include <string>
#include <map>
namespace {
struct Mark { const char* name; cudaEvent_t ev; int launches; };
bool g_prof_on = false;
std::vector<Mark> g_marks;
std::vector<cudaEvent_t> g_pool;
long long g_last_launches = 0;
}  // namespace
long long g_vistaf_launches = 0;

void vistaf_stage_mark(const char* name, cudaStream_t st) {
  if (!g_prof_on) return;
  const int launches_before = (int)(g_vistaf_launches - g_last_launches);
  g_last_launches = g_vistaf_launches;
  cudaEvent_t ev;
  if (!g_pool.empty()) { ev = g_pool.back(); g_pool.pop_back(); }
  else if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st);
  if (!g_marks.empty()) g_marks.back().launches = launches_before;
  g_marks.push_back(Mark{name, ev, 0});
}

extern "C" int vistaf_profile_enable(int on) {
  g_prof_on = on != 0;
  for (auto& m : g_marks) g_pool.push_back(m.ev);
  g_marks.clear();
  return 0;
}

extern "C" long long vistaf_launch_count(void) { return g_vistaf_launches; }

extern "C" int vistaf_profile_read(VistafStageTime* out, int max_stages, int* n_out) {
  if (!out || !n_out) return vistaf_set_err("vistaf_profile_read: null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  *n_out = 0;
  if (g_marks.empty()) return 0;
  CK(cudaEventSynchronize(g_marks.back().ev));
  std::vector<std::string> order;
  std::map<std::string, VistafStageTime> acc;
  for (size_t i = 0; i + 1 < g_marks.size(); i++) {
    const char* nm = g_marks[i].name;
    if (!nm || !nm[0]) continue;                       // "" = boundary mark (end of a call)
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, g_marks[i].ev, g_marks[i + 1].ev));
    auto it = acc.find(nm);
    if (it == acc.end()) {
      VistafStageTime s;
      memset(&s, 0, sizeof(s));
      snprintf(s.name, sizeof(s.name), "%s", nm);
      it = acc.emplace(nm, s).first;
      order.push_back(nm);
    }
    it->second.ms += ms;
    it->second.calls += 1;
    it->second.launches += g_marks[i].launches;
  }
  for (auto& m : g_marks) g_pool.push_back(m.ev);
  g_marks.clear();
  int n = 0;
  for (auto& k : order) { if (n >= max_stages) break; out[n++] = acc[k]; }
  *n_out = n;
  return 0;
}
