// Context, filterbank and error plumbing of the C ABI (include/yoho_b200.h).
#include "common.cuh"

#include <dlfcn.h>

#include <atomic>
#include <cmath>
#include <mutex>
#include <cstdlib>
#include <cstring>
#include <new>
#include <algorithm>
#include <vector>

namespace yoho {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
  return code;
}

bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("YOHO_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

// ---- launch counter + event profiler ---------------------------------------------------------
static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

struct ProfRecord { int cls; double work; cudaEvent_t start, stop; };
static std::mutex g_prof_mutex;
static bool g_prof_on = false;
static std::vector<ProfRecord> g_prof_records;
static std::vector<cudaEvent_t> g_event_pool;

static cudaEvent_t pool_event() {
  if (!g_event_pool.empty()) { cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

ProfScope::ProfScope(cudaStream_t s, int cls, double work) : stream(s), slot(-1) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  if (!g_prof_on) return;
  ProfRecord r{cls, work, pool_event(), pool_event()};
  if (!r.start || !r.stop) return;
  cudaEventRecord(r.start, s);
  g_prof_records.push_back(r);
  slot = int(g_prof_records.size()) - 1;
}
ProfScope::~ProfScope() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  if (slot < int(g_prof_records.size())) cudaEventRecord(g_prof_records[slot].stop, stream);
}

// Setup calls allocate on a given device: switch to it and put the caller's device back afterwards.
struct DeviceGuard {
  int prev = -1;
  cudaError_t error = cudaSuccess;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != device) error = cudaSetDevice(device); else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace yoho

using namespace yoho;

extern "C" long long yoho_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int yoho_prof_enable(int on) {
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  for (auto& r : g_prof_records) { g_event_pool.push_back(r.start); g_event_pool.push_back(r.stop); }
  g_prof_records.clear();
  g_prof_on = on != 0;
  return YOHO_OK;
}

extern "C" int yoho_prof_num_classes(void) { return kProfNumClasses; }

extern "C" const char* yoho_prof_class_name(int cls) {
  static const char* names[kProfNumClasses] = {"logmel", "normalize", "gemm_tcgen05", "attention_tcgen05", "layernorm", "other", "decoder_skinny_gemm", "decoder_cache_attention"};
  return (cls >= 0 && cls < kProfNumClasses) ? names[cls] : "";
}

extern "C" int yoho_prof_collect(double* ms, double* work, long long* launches, int n) {
  if (!ms || !work || !launches || n < kProfNumClasses) return fail(YOHO_ERR_INVALID, "yoho_prof_collect: need %d slots", kProfNumClasses);
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  for (int i = 0; i < n; ++i) { ms[i] = 0.0; work[i] = 0.0; launches[i] = 0; }
  for (auto& r : g_prof_records) {
    cudaError_t e = cudaEventSynchronize(r.stop);
    if (e != cudaSuccess) return fail(YOHO_ERR_CUDA, "yoho_prof_collect: %s", cudaGetErrorString(e));
    float t = 0.0f;
    e = cudaEventElapsedTime(&t, r.start, r.stop);
    if (e != cudaSuccess) return fail(YOHO_ERR_CUDA, "yoho_prof_collect: %s", cudaGetErrorString(e));
    ms[r.cls] += t; work[r.cls] += r.work; launches[r.cls] += 1;
  }
  return YOHO_OK;
}

extern "C" int yoho_abi_version(void) { return YOHO_ABI_VERSION; }
extern "C" const char* yoho_last_error(void) { return g_error; }

extern "C" int yoho_ctx_create(int device, yoho_ctx** out) {
  if (!out) return fail(YOHO_ERR_INVALID, "yoho_ctx_create: out is null");
  *out = nullptr;
  int count = 0;
  YOHO_CUDA_CHECK(cudaGetDeviceCount(&count));
  if (device < 0 || device >= count) return fail(YOHO_ERR_INVALID, "yoho_ctx_create: device %d of %d", device, count);
  DeviceGuard guard(device);                        // the caller's current device is restored on every return path
  if (guard.error != cudaSuccess) return fail(YOHO_ERR_CUDA, "yoho_ctx_create: cudaSetDevice(%d): %s", device, cudaGetErrorString(guard.error));
  cudaDeviceProp prop;
  YOHO_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(YOHO_ERR_ARCH, "yoho_ctx_create: device %d (%s) is sm_%d%d; this library is built for sm_100a only and has no fallback",
                device, prop.name, prop.major, prop.minor);
  yoho_ctx* ctx = new (std::nothrow) yoho_ctx();
  if (!ctx) return fail(YOHO_ERR_INVALID, "yoho_ctx_create: out of host memory");
  ctx->device = device;
  ctx->num_sms = prop.multiProcessorCount;
  ctx->cc_major = prop.major;
  ctx->cc_minor = prop.minor;

  // Periodic Hann window and twiddles of the 400-point transform go to the frontend's constant tables.
  if (int rc = frontend_upload_constants(&ctx->window_sum)) { delete ctx; return rc; }
  cudaError_t e;
  // cuTensorMapEncodeTiled without linking libcuda.
  cudaDriverEntryPointQueryResult qres;
  void* fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    yoho_ctx_destroy(ctx);
    return fail(YOHO_ERR_CUDA, "yoho_ctx_create: cuTensorMapEncodeTiled entry point unavailable");
  }
  ctx->encode_tiled = fn;
  *out = ctx;
  return YOHO_OK;
}

extern "C" int yoho_ctx_destroy(yoho_ctx* ctx) {
  if (!ctx) return YOHO_OK;
  delete ctx;
  return YOHO_OK;
}

extern "C" int yoho_filterbank_create(yoho_ctx* ctx, const float* host_filters, int n_mels, int n_bins,
                                      yoho_filterbank** out) {
  if (!ctx || !host_filters || !out) return fail(YOHO_ERR_INVALID, "yoho_filterbank_create: null argument");
  *out = nullptr;
  if (n_mels <= 0 || n_bins <= 0) return fail(YOHO_ERR_INVALID, "yoho_filterbank_create: shape (%d, %d)", n_mels, n_bins);
  DeviceGuard guard(ctx->device);
  if (guard.error != cudaSuccess) return fail(YOHO_ERR_CUDA, "yoho_filterbank_create: cudaSetDevice(%d): %s", ctx->device, cudaGetErrorString(guard.error));
  if (n_mels > kMelMaxBands) return fail(YOHO_ERR_UNSUPPORTED, "yoho_filterbank_create: n_mels=%d > %d", n_mels, kMelMaxBands);
  if (n_bins > kMelBins) return fail(YOHO_ERR_UNSUPPORTED, "yoho_filterbank_create: n_bins=%d > %d", n_bins, kMelBins);
  std::vector<int> lo(n_mels, 0), len(n_mels, 0);
  for (int m = 0; m < n_mels; ++m) {
    int first = -1, last = -1;
    for (int k = 0; k < n_bins; ++k)
      if (host_filters[size_t(m) * n_bins + k] != 0.0f) { if (first < 0) first = k; last = k; }
    if (first >= 0) { lo[m] = first; len[m] = last - first + 1; }
  }
  std::vector<int> order(n_mels);
  for (int m = 0; m < n_mels; ++m) order[m] = m;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return len[a] > len[b]; });
  // |2 X[k]|^2 comes out of the packed-pair FFT; stft scales by 1/sum(w) (scaling='spectrum').
  const double scale = 0.25 / (ctx->window_sum * ctx->window_sum);
  // Groups of eight bands (longest first) share one trip count: one warp-round of the kernel's mel phase.  The groups are
  // dealt to the CTA's six warps longest-processing-time first and stored [round][warp]; a warp without a group in a round
  // finds a filler (trip 0).
  const int n_real = (n_mels + kMelGroup - 1) / kMelGroup;
  std::vector<int> trips(n_real);
  for (int g = 0; g < n_real; ++g) {
    trips[g] = (len[order[kMelGroup * g]] + 1) & ~1;              // the group's longest band comes first; walked two bins a step
    if (trips[g] > n_bins - (n_bins & 1)) return fail(YOHO_ERR_UNSUPPORTED, "yoho_filterbank_create: band of %d bins", len[order[kMelGroup * g]]);
  }
  std::vector<std::vector<int>> mine(kMelWarps);
  std::vector<int> load(kMelWarps, 0);
  for (int g = 0; g < n_real; ++g) {                              // trips are already in descending order
    int best = 0;
    for (int wv = 1; wv < kMelWarps; ++wv) if (load[wv] < load[best]) best = wv;
    mine[best].push_back(g);
    load[best] += trips[g] / 2 + 2;                               // steps + the round's fixed cost
  }
  size_t rounds = 0;
  for (const auto& v : mine) rounds = v.size() > rounds ? v.size() : rounds;
  if (rounds > size_t(kMelMaxRounds)) return fail(YOHO_ERR_UNSUPPORTED, "yoho_filterbank_create: %zu rounds of band groups", rounds);
  const int n_groups = n_real;
  std::vector<int4> bands(size_t(kMelGroup) * n_groups + 8, make_int4(-1, -1, -1, -1));   // descriptors, then schedule[32]
  int* sched = reinterpret_cast<int*>(bands.data() + size_t(kMelGroup) * n_groups);
  for (size_t r = 0; r < rounds; ++r)
    for (int wv = 0; wv < kMelWarps; ++wv)
      if (r < mine[wv].size()) sched[r * kMelWarps + wv] = mine[wv][r];
  std::vector<float2> w;
  for (int g = 0; g < n_groups; ++g) {
    const int trip = trips[g];
    int4* dst = bands.data() + size_t(kMelGroup) * g;
    for (int h = 0; h < kMelGroup; ++h) {
      const int idx = kMelGroup * g + h;
      if (idx >= n_mels) { dst[h] = make_int4(-1, 0, int(w.size()) - trip, trip); continue; }   // walks its neighbour's weights, stores nothing
      const int m = order[idx];
      const int first = lo[m] + trip <= n_bins ? lo[m] : n_bins - trip;   // keep the walk inside the power tile
      dst[h] = make_int4(m, first, int(w.size()), trip);
      for (int jj = 0; jj < trip; ++jj) {
        const int k = first + jj;
        const float v = (k >= lo[m] && k < lo[m] + len[m]) ? float(double(host_filters[size_t(m) * n_bins + k]) * scale) : 0.0f;
        w.push_back(make_float2(v, v));
      }
    }
  }
  if (w.empty()) w.push_back(make_float2(0.0f, 0.0f));
  if (w.size() > size_t(kMelMaxWeights))
    return fail(YOHO_ERR_UNSUPPORTED, "yoho_filterbank_create: %zu walked bins exceed the kernel's %d-entry weight tile", w.size(), kMelMaxWeights);
  yoho_filterbank* fb = new (std::nothrow) yoho_filterbank();
  if (!fb) return fail(YOHO_ERR_INVALID, "yoho_filterbank_create: out of host memory");
  fb->device = ctx->device;
  fb->n_mels = n_mels;
  fb->n_bins = n_bins;
  fb->n_groups = n_groups;
  fb->n_rounds = int(rounds);
  fb->n_weights = int(w.size());
  fb->power_scale = scale;
  cudaError_t e;
  if ((e = cudaMalloc(&fb->d_bands, bands.size() * sizeof(int4))) != cudaSuccess ||
      (e = cudaMalloc(&fb->d_weights, w.size() * sizeof(float2))) != cudaSuccess ||
      (e = cudaMemcpy(fb->d_bands, bands.data(), bands.size() * sizeof(int4), cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(fb->d_weights, w.data(), w.size() * sizeof(float2), cudaMemcpyHostToDevice)) != cudaSuccess) {
    yoho_filterbank_destroy(fb);
    return fail(YOHO_ERR_CUDA, "yoho_filterbank_create: %s", cudaGetErrorString(e));
  }
  *out = fb;
  return YOHO_OK;
}

extern "C" int yoho_filterbank_destroy(yoho_filterbank* fb) {
  if (!fb) return YOHO_OK;
  if (fb->d_bands) cudaFree(fb->d_bands);
  if (fb->d_weights) cudaFree(fb->d_weights);
  delete fb;
  return YOHO_OK;
}

// ---- optional gather of encoder states -------------------------------------------------------------
// NCCL is not linked: the symbol is taken from the libnccl the host process already uses (torch's, or the JAX
// runtime's), so the communicator handed in and the function called belong to the same library instance.
extern "C" int yoho_allgather_states(yoho_ctx* ctx, void* nccl_comm, const void* states, void* gathered,
                                     size_t elems_per_rank, int dtype, yoho_stream_t stream) {
  if (!ctx || !nccl_comm || !states || !gathered) return fail(YOHO_ERR_INVALID, "yoho_allgather_states: null argument");
  if (dtype != YOHO_DTYPE_F32 && dtype != YOHO_DTYPE_BF16) return fail(YOHO_ERR_INVALID, "yoho_allgather_states: dtype=%d", dtype);
  using AllGatherFn = int (*)(const void*, void*, size_t, int, void*, cudaStream_t);
  static AllGatherFn fn = nullptr;
  static std::mutex mu;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!fn) {
      void* sym = dlsym(RTLD_DEFAULT, "ncclAllGather");
      for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
        if (sym) break;
        void* h = dlopen(name, RTLD_LAZY | RTLD_NOLOAD);
        if (!h) h = dlopen(name, RTLD_LAZY);
        if (h) sym = dlsym(h, "ncclAllGather");
      }
      if (!sym) return fail(YOHO_ERR_UNSUPPORTED, "yoho_allgather_states: no ncclAllGather in this process (load libnccl.so.2 first)");
      fn = reinterpret_cast<AllGatherFn>(sym);
    }
  }
  const int nccl_dtype = dtype == YOHO_DTYPE_F32 ? 7 /* ncclFloat32 */ : 9 /* ncclBfloat16 */;
  const int rc = fn(states, gathered, elems_per_rank, nccl_dtype, nccl_comm, static_cast<cudaStream_t>(stream));
  if (rc != 0) return fail(YOHO_ERR_CUDA, "yoho_allgather_states: ncclAllGather returned %d", rc);
  return YOHO_OK;
}
