// Optional per-kernel timing for bench.py's roofline line: when enabled, the dominant kernel of each
// entry point is bracketed by CUDA events ON THE STREAM IT IS LAUNCHED ON; dlas_prof_read()
// synchronises those events and returns the summed duration and the launch count per slot.
#include "common.cuh"
#include <vector>
#include <mutex>

namespace dlas {
namespace {
struct Pair { cudaEvent_t a, b; int slot; };
std::mutex g_mu;
std::vector<Pair> g_pairs;
bool g_on = false;
}  // namespace

bool prof_on() { return g_on; }

// returns an index to pass to prof_end, or -1
int prof_begin(int slot, cudaStream_t st) {
  if (!g_on) return -1;
  Pair p;
  p.slot = slot;
  if (cudaEventCreate(&p.a) != cudaSuccess || cudaEventCreate(&p.b) != cudaSuccess) return -1;
  cudaEventRecord(p.a, st);
  std::lock_guard<std::mutex> lk(g_mu);
  g_pairs.push_back(p);
  return (int)g_pairs.size() - 1;
}

void prof_end(int idx, cudaStream_t st) {
  if (idx < 0) return;
  std::lock_guard<std::mutex> lk(g_mu);
  cudaEventRecord(g_pairs[idx].b, st);
}
}  // namespace dlas

extern "C" int dlas_prof_enable(int on) {
  dlas::g_on = on != 0;
  return DLAS_OK;
}

extern "C" int dlas_prof_read(int slot, float* ms_sum_host, int* count_host) {
  using namespace dlas;
  std::lock_guard<std::mutex> lk(g_mu);
  float sum = 0.0f;
  int cnt = 0;
  std::vector<Pair> keep;
  for (auto& p : g_pairs) {
    if (p.slot != slot) { keep.push_back(p); continue; }
    float ms = 0.0f;
    DLAS_CUDA(cudaEventSynchronize(p.b));
    DLAS_CUDA(cudaEventElapsedTime(&ms, p.a, p.b));
    sum += ms;
    ++cnt;
    cudaEventDestroy(p.a);
    cudaEventDestroy(p.b);
  }
  g_pairs.swap(keep);
  if (ms_sum_host) *ms_sum_host = sum;
  if (count_host) *count_host = cnt;
  return DLAS_OK;
}
