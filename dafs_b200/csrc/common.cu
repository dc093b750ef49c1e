// Lifecycle + error plumbing of libdafs_cuda, and small caches of pinned host
// buffers and large device buffers so repeated calls do not pay cudaHostAlloc /
// cudaMalloc / cudaFree every time.
#include "common.cuh"

#include <algorithm>
#include <cstdlib>
#include <map>
#include <cstring>
#include <mutex>
#include <unordered_set>
#include <vector>

namespace dafs {

// dafs_cuda_progressive keeps many small launches in flight on their own streams next to
// launches that run for tens of milliseconds.  With the default of 8 hardware work queues,
// streams share queues and a short launch can sit behind a long one of another stream; ask
// for the maximum (only effective if the CUDA context does not exist yet; never overrides
// a value the user has set).
namespace {
struct MaxConnections {
  MaxConnections() { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0); }
} g_max_connections;
}  // namespace

thread_local DDPackLend g_dd_pack_lend = {nullptr, 0};
thread_local cudaStream_t g_dd_stream = nullptr, g_dd_poll_stream = nullptr;

std::string& last_error() {
  static thread_local std::string e;
  return e;
}

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  last_error() = buf;
  return code;
}

namespace {
std::mutex g_pin_mu;
std::multimap<size_t, void*> g_pin_free;   // capacity -> buffer
std::map<void*, size_t> g_pin_cap;
size_t g_pin_cached = 0;
const size_t kPinCacheLimit = size_t(6) << 30;
}  // namespace

void* pinned_alloc(size_t bytes) {
  if (bytes == 0) bytes = 1;
  {
    std::lock_guard<std::mutex> g(g_pin_mu);
    auto it = g_pin_free.lower_bound(bytes);
    if (it != g_pin_free.end() && it->first <= 2 * bytes + (1 << 20)) {
      void* p = it->second;
      g_pin_cached -= it->first;
      g_pin_free.erase(it);
      return p;
    }
  }
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  std::lock_guard<std::mutex> g(g_pin_mu);
  g_pin_cap[p] = bytes;
  return p;
}

void pinned_free(void* p) {
  if (!p) return;
  std::lock_guard<std::mutex> g(g_pin_mu);
  const size_t cap = g_pin_cap[p];
  if (g_pin_cached + cap <= kPinCacheLimit) {
    g_pin_free.emplace(cap, p);
    g_pin_cached += cap;
  } else {
    g_pin_cap.erase(p);
    cudaFreeHost(p);
  }
}

void pinned_trim() {
  std::lock_guard<std::mutex> g(g_pin_mu);
  for (auto& kv : g_pin_free) { g_pin_cap.erase(kv.second); cudaFreeHost(kv.second); }
  g_pin_free.clear();
  g_pin_cached = 0;
}

// ---- device buffers ------------------------------------------------------------
// cudaFree of a multi-GB workspace costs tens of milliseconds (and synchronises
// the device); plans are created and destroyed per call by the reference-facing
// adapters, so large device buffers are parked here instead and handed to the
// next plan that asks for a similar size on the same device.
namespace {
struct DevBlock { void* p; size_t cap; int dev; };
std::mutex g_dev_mu;
std::vector<DevBlock> g_dev_free;
std::map<void*, std::pair<size_t, int>> g_dev_cap;   // cached-class allocations in use
size_t g_dev_cached = 0;
const size_t kDevCacheMin = 256;                      // smaller requests are rounded up: a cudaFree of even a 4-byte
                                                      // counter waits for every kernel running on the device
const size_t kDevSmall = size_t(1) << 20;
const size_t kDevCacheLimit = size_t(32) << 30;
}  // namespace

cudaError_t dev_malloc(void** out, size_t bytes) {
  *out = nullptr;
  if (bytes < kDevCacheMin) bytes = kDevCacheMin;
  int dev = 0;
  cudaGetDevice(&dev);
  if (bytes >= kDevCacheMin) {
    std::lock_guard<std::mutex> g(g_dev_mu);
    int best = -1;
    for (int k = 0; k < (int)g_dev_free.size(); ++k) {
      const DevBlock& b = g_dev_free[k];
      const size_t most = bytes < kDevSmall ? std::max<size_t>(2 * bytes, 4096) : bytes + bytes / 2;
      if (b.dev == dev && b.cap >= bytes && b.cap <= most && (best < 0 || b.cap < g_dev_free[best].cap)) best = k;
    }
    if (best >= 0) {
      const DevBlock b = g_dev_free[best];
      g_dev_free.erase(g_dev_free.begin() + best);
      g_dev_cached -= b.cap;
      g_dev_cap[b.p] = {b.cap, b.dev};
      *out = b.p;
      return cudaSuccess;
    }
  }
  cudaError_t e = cudaMalloc(out, bytes);
  if (e == cudaErrorMemoryAllocation) {        // give the parked buffers back and try once more
    cudaGetLastError();
    dev_trim();
    e = cudaMalloc(out, bytes);
  }
  if (e == cudaSuccess && bytes >= kDevCacheMin) {
    std::lock_guard<std::mutex> g(g_dev_mu);
    g_dev_cap[*out] = {bytes, dev};
  }
  return e;
}

void dev_free(void* p) {
  if (!p) return;
  {
    std::lock_guard<std::mutex> g(g_dev_mu);
    auto it = g_dev_cap.find(p);
    if (it != g_dev_cap.end()) {
      const size_t cap = it->second.first;
      const int dev = it->second.second;
      g_dev_cap.erase(it);
      if (g_dev_cached + cap <= kDevCacheLimit) {
        g_dev_free.push_back({p, cap, dev});
        g_dev_cached += cap;
        return;
      }
    }
  }
  cudaFree(p);
}

void dev_trim() {
  std::vector<DevBlock> blocks;
  {
    std::lock_guard<std::mutex> g(g_dev_mu);
    blocks.swap(g_dev_free);
    g_dev_cached = 0;
  }
  int cur = 0;
  cudaGetDevice(&cur);
  for (const DevBlock& b : blocks) { cudaSetDevice(b.dev); cudaFree(b.p); }
  cudaSetDevice(cur);
}

cudaError_t dev_info(int device, DevInfo* out) {
  static std::mutex mu;
  static DevInfo cache[64];
  static bool have[64] = {false};
  std::lock_guard<std::mutex> g(mu);
  const int k = device & 63;
  if (!have[k]) {
    int sms = 0, smem = 0;
    cudaError_t e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e != cudaSuccess) return e;
    cache[k].sms = sms; cache[k].smem_optin = (size_t)smem;
    have[k] = true;
  }
  *out = cache[k];
  return cudaSuccess;
}

// ---- device-resident batches ------------------------------------------------------------
namespace {
struct DeviceBatchOwner { int device; void* pinned; void* d_row_ptr; void* d_col; void* d_val; };
std::mutex g_devbatch_mu;
std::unordered_set<void*> g_devbatch;
}  // namespace

int make_device_batch(int64_t units, int64_t nnz, const int64_t* h_row_off, const int64_t* d_nz_off,
                      int32_t* d_row_ptr, uint32_t* d_col, float* d_val, cudaStream_t st, dafs_csr_batch* out) {
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t b_off = sizeof(int64_t) * (units + 1);
  char* pin = (char*)pinned_alloc(2 * up(b_off) + 256);
  if (!pin) return fail(DAFS_CUDA_ENOMEM, "pinned allocation failed");
  int64_t* row_off = (int64_t*)pin;
  int64_t* nz_off = (int64_t*)(pin + up(b_off));
  std::memcpy(row_off, h_row_off, b_off);
  cudaError_t e = cudaMemcpyAsync(nz_off, d_nz_off, b_off, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { pinned_free(pin); return fail(DAFS_CUDA_ECUDA, "copy of nz_off failed: %s", cudaGetErrorString(e)); }
  DeviceBatchOwner* o = new DeviceBatchOwner();
  cudaGetDevice(&o->device);
  o->pinned = pin; o->d_row_ptr = d_row_ptr; o->d_col = d_col; o->d_val = d_val;
  { std::lock_guard<std::mutex> g(g_devbatch_mu); g_devbatch.insert(o); }
  out->n_units = units; out->nnz = nnz; out->row_off = row_off; out->nz_off = nz_off;
  out->row_ptr = d_row_ptr; out->col = d_col; out->val = d_val; out->owner_ = o;
  return DAFS_CUDA_OK;
}

bool is_device_batch(const dafs_csr_batch* b) {
  if (!b || !b->owner_) return false;
  std::lock_guard<std::mutex> g(g_devbatch_mu);
  return g_devbatch.count(b->owner_) != 0;
}

bool free_device_batch(dafs_csr_batch* b) {
  if (!b || !b->owner_) return false;
  DeviceBatchOwner* o = nullptr;
  {
    std::lock_guard<std::mutex> g(g_devbatch_mu);
    auto it = g_devbatch.find(b->owner_);
    if (it == g_devbatch.end()) return false;
    o = static_cast<DeviceBatchOwner*>(*it);
    g_devbatch.erase(it);
  }
  int cur = 0;
  cudaGetDevice(&cur);
  if (cur != o->device) cudaSetDevice(o->device);
  dev_free(o->d_row_ptr); dev_free(o->d_col); dev_free(o->d_val);
  if (cur != o->device) cudaSetDevice(cur);
  pinned_free(o->pinned);
  delete o;
  return true;
}

// FP32 issue-rate probe: 8 independent chains per thread, 4 CTAs of 256 per SM.
// FMA = true counts 2 flop per FFMA; false issues FMUL + FADD (what the bit-exact
// pair-HMM arithmetic is restricted to, pairhmm_math.cuh) and counts 1 each.
template <bool FMA>
__global__ void __launch_bounds__(256) fp32_probe_kernel(float* out, float a, float b, int iters) {
  float v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = (float)(threadIdx.x + k) * 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = FMA ? __fmaf_rn(v[k], a, b) : __fadd_rn(__fmul_rn(v[k], a), b);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 12345.678f) out[0] = s;
}

// FP64 issue-rate probe (DFMA, 2 flop each), same shape: the denominator of the McCaskill roofline.
__global__ void __launch_bounds__(256) fp64_probe_kernel(double* out, double a, double b, int iters) {
  double v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = (double)(threadIdx.x + k) * 1e-3;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __fma_rn(v[k], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 12345.678) out[0] = s;
}

}  // namespace dafs

extern "C" {

int dafs_cuda_fp64_probe(double* fma_tflops) {
  int dev = 0, sms = 0;
  DAFS_CUDA_TRY(cudaGetDevice(&dev));
  DAFS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  double* out = nullptr;
  DAFS_CUDA_TRY(cudaMalloc(&out, 8));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 1 << 13, grid = sms * 8;
  double best = 0;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    dafs::fp64_probe_kernel<<<grid, 256>>>(out, 0.999, 1e-3, iters);
    cudaEventRecord(e1);
    DAFS_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double tf = 2.0 * 8 * iters * 256.0 * grid / (ms * 1e-3) / 1e12;
    if (rep && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(out);
  DAFS_CUDA_TRY(cudaGetLastError());
  if (fma_tflops) *fma_tflops = best;
  return DAFS_CUDA_OK;
}

int dafs_cuda_fp32_probe(double* fma_tflops, double* muladd_tflops) {
  int dev = 0, sms = 0;
  DAFS_CUDA_TRY(cudaGetDevice(&dev));
  DAFS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  float* out = nullptr;
  DAFS_CUDA_TRY(cudaMalloc(&out, 4));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 1 << 15, grid = sms * 8;
  double best[2] = {0, 0};
  for (int mode = 0; mode < 2; ++mode)
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) dafs::fp32_probe_kernel<true><<<grid, 256>>>(out, 0.999f, 1e-3f, iters);
      else dafs::fp32_probe_kernel<false><<<grid, 256>>>(out, 0.999f, 1e-3f, iters);
      cudaEventRecord(e1);
      DAFS_CUDA_TRY(cudaEventSynchronize(e1));
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double tf = 2.0 * 8 * iters * 256.0 * grid / (ms * 1e-3) / 1e12;
      if (rep && tf > best[mode]) best[mode] = tf;
    }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(out);
  DAFS_CUDA_TRY(cudaGetLastError());
  if (fma_tflops) *fma_tflops = best[0];
  if (muladd_tflops) *muladd_tflops = best[1];
  return DAFS_CUDA_OK;
}

const char* dafs_cuda_last_error(void) { return dafs::last_error().c_str(); }

int dafs_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int dafs_cuda_init(int ngpu) {
  const int n = dafs_cuda_device_count();
  if (n <= 0) return dafs::fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  if (ngpu < 0 || ngpu > n) return dafs::fail(DAFS_CUDA_ENODEV, "requested %d GPUs, %d visible", ngpu, n);
  const int use = ngpu == 0 ? n : ngpu;
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < use; ++d) {
    cudaDeviceProp p;
    DAFS_CUDA_TRY(cudaGetDeviceProperties(&p, d));
    if (p.major != 10)
      return dafs::fail(DAFS_CUDA_ENODEV, "device %d is sm_%d%d; this library is built for sm_100a only", d, p.major, p.minor);
    DAFS_CUDA_TRY(cudaSetDevice(d));
    DAFS_CUDA_TRY(cudaFree(0));
  }
  cudaSetDevice(cur);
  return DAFS_CUDA_OK;
}

void dafs_cuda_shutdown(void) { dafs::pinned_trim(); dafs::dev_trim(); }

void dafs_cuda_free_batch(dafs_csr_batch* b) {
  if (!b) return;
  if (!dafs::free_device_batch(b)) dafs::pinned_free((void*)b->owner_);
  *b = dafs_csr_batch{};
}

int dafs_cuda_batch_is_device(const dafs_csr_batch* b) { return b && dafs::is_device_batch(b) ? 1 : 0; }

int dafs_cuda_batch_to_host(const dafs_csr_batch* b, dafs_csr_batch* out) {
  if (!b || !out) return dafs::fail(DAFS_CUDA_EINVAL, "NULL argument");
  *out = dafs_csr_batch{};
  const int64_t units = b->n_units, rows = units >= 0 && b->row_off ? b->row_off[units] : 0, nnz = b->nnz;
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t b_off = sizeof(int64_t) * (units + 1);
  char* base = (char*)dafs::pinned_alloc(2 * up(b_off) + up(sizeof(int32_t) * rows) + up(sizeof(uint32_t) * nnz) + up(sizeof(float) * nnz) + 256);
  if (!base) return dafs::fail(DAFS_CUDA_ENOMEM, "pinned allocation failed");
  int64_t* row_off = (int64_t*)base;
  int64_t* nz_off = (int64_t*)(base + up(b_off));
  int32_t* row_ptr = (int32_t*)((char*)nz_off + up(b_off));
  uint32_t* col = (uint32_t*)((char*)row_ptr + up(sizeof(int32_t) * rows));
  float* val = (float*)((char*)col + up(sizeof(uint32_t) * nnz));
  std::memcpy(row_off, b->row_off, b_off);
  std::memcpy(nz_off, b->nz_off, b_off);
  cudaError_t e = rows ? cudaMemcpy(row_ptr, b->row_ptr, sizeof(int32_t) * rows, cudaMemcpyDefault) : cudaSuccess;
  if (e == cudaSuccess && nnz) e = cudaMemcpy(col, b->col, sizeof(uint32_t) * nnz, cudaMemcpyDefault);
  if (e == cudaSuccess && nnz) e = cudaMemcpy(val, b->val, sizeof(float) * nnz, cudaMemcpyDefault);
  if (e != cudaSuccess) { dafs::pinned_free(base); return dafs::fail(DAFS_CUDA_ECUDA, "copy of the batch failed: %s", cudaGetErrorString(e)); }
  out->n_units = units; out->nnz = nnz; out->row_off = row_off; out->nz_off = nz_off; out->row_ptr = row_ptr; out->col = col; out->val = val;
  out->owner_ = base;
  return DAFS_CUDA_OK;
}

}  // extern "C"
