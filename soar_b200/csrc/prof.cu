// Launch accounting and optional per-kernel CUDA-event timing for bench.py:
//   * soar_launch_count()  — how many of OUR kernels were launched by this process
//     (bench.py reports the difference across its timed region as `gpu_launches`);
//   * soar_prof_enable(1)  — from then on every launch is bracketed by two events
//     recorded on the launching stream; soar_prof_summary() synchronises and
//     returns, per kernel id, the launch count and the summed device time. This is
//     what bench.py's `roofline.achieved` is computed from (events on the stream
//     the kernel is launched on, never a profiler).
#include <atomic>
#include <mutex>
#include <vector>
#include "common.cuh"

namespace soar {

static std::atomic<unsigned long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
struct ProfRec {
  int id;
  cudaEvent_t e0, e1;
};
static std::vector<ProfRec> g_recs;
static std::vector<cudaEvent_t> g_pool;

static cudaEvent_t get_event() {
  if (!g_pool.empty()) {
    cudaEvent_t e = g_pool.back();
    g_pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

void KernelScope::count_extra_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

KernelScope::KernelScope(int id, cudaStream_t stream) : id_(id), stream_(stream), e1_(nullptr) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (g_prof_on.load(std::memory_order_relaxed)) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    cudaEvent_t e0 = get_event();
    e1_ = get_event();
    cudaEventRecord(e0, stream_);
    g_recs.push_back({id_, e0, (cudaEvent_t)e1_});
  }
}

KernelScope::~KernelScope() {
  if (e1_) cudaEventRecord((cudaEvent_t)e1_, stream_);
}

static const char* kKernelNames[K_COUNT] = {
    "preprocess_fwd", "scan_lookback", "duplicate_keys", "radix_hist", "radix_hist_scan",
    "radix_onesweep", "tile_ranges", "blend_fwd", "blend_bwd", "geom_bwd", "knn_bbox",
    "knn_morton", "knn_gather", "knn_boxes", "knn_search", "lbs_fwd", "lbs_bwd", "lbs_query",
    "smplx_shape", "smplx_posedirs", "smplx_joints", "smplx_chain", "smplx_skin", "post_fwd",
    "post_bwd", "lbs_query_build", "adam_step", "densify_stats", "densify_select",
    "mask_index", "rows_gather", "split_rows", "occgrid_splat", "contour_padding", "lbs_query_keys",
    "lbs_query_tail"};

}  // namespace soar

using namespace soar;

extern "C" {

unsigned long long soar_launch_count(void) { return g_launches.load(); }

void soar_prof_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& r : g_recs) {
    g_pool.push_back(r.e0);
    g_pool.push_back(r.e1);
  }
  g_recs.clear();
  g_prof_on.store(on ? 1 : 0);
}

int soar_prof_summary(int max_ids, int* counts, float* total_ms) {
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (int i = 0; i < max_ids; ++i) {
    counts[i] = 0;
    total_ms[i] = 0.f;
  }
  for (auto& r : g_recs) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) != cudaSuccess) continue;
    if (r.id >= 0 && r.id < max_ids) {
      counts[r.id] += 1;
      total_ms[r.id] += ms;
    }
  }
  return K_COUNT;
}

const char* soar_kernel_name(int id) { return (id >= 0 && id < K_COUNT) ? kKernelNames[id] : ""; }

}  // extern "C"
