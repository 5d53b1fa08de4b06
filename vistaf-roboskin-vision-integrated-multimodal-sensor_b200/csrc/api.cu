// Error plumbing, launch accounting, stage timers and the CUDA-graph replay cache of the C ABI.
#include "../../include/vistaf.h"
#include "common.cuh"
#include <string.h>
#include <stdlib.h>
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

// last error of the CALLING THREAD (vistaf_last_error), so concurrent callers do not clobber each other's message
static thread_local char t_vistaf_err[512] = "";

int vistaf_set_err(const char* what, cudaError_t e, const char* file, int line) {
  snprintf(t_vistaf_err, sizeof(t_vistaf_err), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  return 1;
}

extern "C" const char* vistaf_last_error(void) { return t_vistaf_err; }
extern "C" int vistaf_version(void) { return 200; }

std::atomic<long long> g_vistaf_launches{0};
extern "C" long long vistaf_launch_count(void) { return g_vistaf_launches.load(std::memory_order_relaxed); }

// ------------------------------------------------------------------ stage timers (CUDA events on the launch stream)
// Off by default (no events recorded).  When enabled, every vistaf_stage_mark(name, stream) records an event; the time
// between two consecutive marks is charged to the earlier mark's name.  bench.py reads the totals to report which
// stage dominates and its achieved bandwidth, measured live on the stream the kernels run on.  One mutex guards the
// mark list: marks from concurrent threads interleave (meaningless as timings) but never corrupt the list.
namespace {
struct Mark { const char* name; cudaEvent_t ev; int launches; };
std::mutex g_prof_mu;
std::atomic<bool> g_prof_on{false};
std::vector<Mark> g_marks;
std::vector<cudaEvent_t> g_pool;
long long g_last_launches = 0;
}  // namespace

bool vistaf_profiling(void) { return g_prof_on.load(std::memory_order_relaxed); }

void vistaf_stage_mark(const char* name, cudaStream_t st) {
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  const long long now = g_vistaf_launches.load(std::memory_order_relaxed);
  const int launches_before = (int)(now - g_last_launches);
  g_last_launches = now;
  cudaEvent_t ev;
  if (!g_pool.empty()) { ev = g_pool.back(); g_pool.pop_back(); }
  else if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st);
  if (!g_marks.empty()) g_marks.back().launches = launches_before;
  g_marks.push_back(Mark{name, ev, 0});
}

extern "C" int vistaf_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on.store(on != 0);
  for (auto& m : g_marks) g_pool.push_back(m.ev);
  g_marks.clear();
  g_last_launches = g_vistaf_launches.load(std::memory_order_relaxed);
  return 0;
}

extern "C" int vistaf_profile_read(VistafStageTime* out, int max_stages, int* n_out) {
  if (!out || !n_out) return vistaf_set_err("vistaf_profile_read: null argument", cudaErrorInvalidValue, __FILE__, __LINE__);
  std::lock_guard<std::mutex> lk(g_prof_mu);
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

// ------------------------------------------------------------------ CUDA-graph replay of the batch entry points
// A batch call of the C ABI is a fixed sequence of kernel launches whose arguments depend only on the call's own
// arguments (plan, device pointers, batch size): every data-dependent decision is taken on the device.  The first
// call with a given argument tuple runs eagerly (one-time attribute setup happens there), the second is captured from
// the caller's stream into a graph, instantiated and launched, every later one is a single cudaGraphLaunch -- the
// host enqueues one node list instead of ~1000 kernels.  Anything that fails to capture falls back to eager launches
// for good.  Off while the stage timers are on (they need events between the kernels) and with VISTAF_GRAPHS=0.
namespace {
struct GraphEntry { int calls = 0; bool broken = false; cudaGraphExec_t exec = nullptr; long long nodes = 0; };
std::mutex g_graph_mu;
std::map<VistafGraphKey, GraphEntry> g_graphs;
int g_graphs_on = -1;
long long g_graph_replays = 0;
bool graphs_enabled() {
  if (g_graphs_on < 0) { const char* e = getenv("VISTAF_GRAPHS"); g_graphs_on = (e && atoi(e) == 0) ? 0 : 1; }
  return g_graphs_on != 0 && !vistaf_profiling();
}
}  // namespace

extern "C" int vistaf_set_graphs(int on) { std::lock_guard<std::mutex> lk(g_graph_mu); g_graphs_on = on ? 1 : 0; return 0; }
extern "C" long long vistaf_graph_replays(void) { return g_graph_replays; }

void vistaf_graph_drop(const void* plan) {
  std::lock_guard<std::mutex> lk(g_graph_mu);
  for (auto it = g_graphs.begin(); it != g_graphs.end();) {
    if (it->first.v[1] == (unsigned long long)(uintptr_t)plan) {
      if (it->second.exec) cudaGraphExecDestroy(it->second.exec);
      it = g_graphs.erase(it);
    } else ++it;
  }
}

int vistaf_graph_run(const VistafGraphKey& key, cudaStream_t st, const std::function<int(cudaStream_t)>& body) {
  if (!graphs_enabled()) return body(st);
  std::unique_lock<std::mutex> lk(g_graph_mu);
  if (g_graphs.size() >= 512 && g_graphs.find(key) == g_graphs.end()) { lk.unlock(); return body(st); }   // callers that never repeat a call
  GraphEntry& e = g_graphs[key];
  if (e.broken) { lk.unlock(); return body(st); }
  if (e.exec) {
    if (cudaGraphLaunch(e.exec, st) != cudaSuccess) return vistaf_set_err("cudaGraphLaunch", cudaGetLastError(), __FILE__, __LINE__);
    g_vistaf_launches.fetch_add(e.nodes, std::memory_order_relaxed);
    g_graph_replays++;
    return 0;
  }
  if (e.calls++ == 0) { lk.unlock(); return body(st); }         // first call: eager
  // second call: capture
  if (cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
    (void)cudaGetLastError();
    e.broken = true;
    lk.unlock();
    return body(st);
  }
  const long long l0 = g_vistaf_launches.load(std::memory_order_relaxed);
  const int rc = body(st);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(st, &graph);
  const long long nodes = g_vistaf_launches.load(std::memory_order_relaxed) - l0;
  if (rc != 0 || ce != cudaSuccess || !graph) {
    (void)cudaGetLastError();
    if (graph) cudaGraphDestroy(graph);
    e.broken = true;
    g_vistaf_launches.fetch_sub(nodes, std::memory_order_relaxed);
    lk.unlock();
    return body(st);                                             // nothing was enqueued by the failed capture
  }
  cudaGraphExec_t exec = nullptr;
  if (cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess || !exec) {
    (void)cudaGetLastError();
    cudaGraphDestroy(graph);
    e.broken = true;
    g_vistaf_launches.fetch_sub(nodes, std::memory_order_relaxed);
    lk.unlock();
    return body(st);
  }
  cudaGraphDestroy(graph);
  e.exec = exec;
  e.nodes = nodes;
  if (cudaGraphLaunch(exec, st) != cudaSuccess) return vistaf_set_err("cudaGraphLaunch", cudaGetLastError(), __FILE__, __LINE__);
  g_graph_replays++;
  return 0;
}
