// Shared host/device helpers for libdafs_cuda (sm_100a only).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/dafs_cuda.h"

namespace dafs {

// Thread-local last-error text behind dafs_cuda_last_error().
std::string& last_error();
int fail(int code, const char* fmt, ...);

#define DAFS_CUDA_TRY(expr)                                                        \
  do {                                                                             \
    cudaError_t e_ = (expr);                                                       \
    if (e_ != cudaSuccess)                                                         \
      return ::dafs::fail(e_ == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM       \
                                                          : DAFS_CUDA_ECUDA,       \
                          "%s:%d: %s -> %s", __FILE__, __LINE__, #expr,            \
                          cudaGetErrorString(e_));                                 \
  } while (0)

// Cached pinned host memory (cudaHostAlloc is slow; results are fetched into it).
void* pinned_alloc(size_t bytes);
void pinned_free(void* p);
void pinned_trim();

// Device memory with a cache: drop-in for cudaMalloc / cudaFree; freed blocks are
// parked and reused by size on the same device until
// dafs_cuda_shutdown() (or an out-of-memory condition) releases them.  Kernels that
// used a parked buffer have always been synchronised by the plan that owned it.
cudaError_t dev_malloc(void** out, size_t bytes);
template <class T> inline cudaError_t dev_malloc(T** out, size_t bytes) { return dev_malloc(reinterpret_cast<void**>(out), bytes); }
void dev_free(void* p);
void dev_trim();

// Device-resident batches: a dafs_csr_batch whose row_off / nz_off are host arrays (pinned) and whose
// row_ptr / col / val live in HBM.  make_device_batch takes ownership of the three device arrays
// (dev_malloc'ed) and copies nz_off from the device; dafs_cuda_free_batch releases either kind.
int make_device_batch(int64_t units, int64_t nnz, const int64_t* h_row_off, const int64_t* d_nz_off,
                      int32_t* d_row_ptr, uint32_t* d_col, float* d_val, cudaStream_t st, dafs_csr_batch* out);
bool is_device_batch(const dafs_csr_batch* b);
bool free_device_batch(dafs_csr_batch* b);   // false: not a device batch

// Streams a caller inside the library lends to the next DD plans of this thread (a plan
// normally makes its own): dafs_cuda_progressive keeps a small pool and reuses an idle stream
// for every batch instead of creating two streams per batch.
extern thread_local cudaStream_t g_dd_stream, g_dd_poll_stream;
// Likewise a device block for the scratch of the pack kernels of dafs_cuda_dd_plan_create_device (dd_pack.cuh):
// the progressive stage owns one grow-only block for all its batches, so that no cudaMalloc / cudaFree (both wait
// for the DD batches in flight) happens per batch.  dd_pack_bytes: what a batch needs.
struct DDPackLend { char* base; size_t cap; };
extern thread_local DDPackLend g_dd_pack_lend;
size_t dd_pack_bytes(int nprob, const dafs_dd_problem* prob);

// SM count and opt-in shared memory per block of a device, queried once
// (cudaGetDeviceProperties costs about a millisecond per call, per-call plans pay it).
struct DevInfo { int sms; size_t smem_optin; };
cudaError_t dev_info(int device, DevInfo* out);

// Symbol classes of the ProbCons alphabet "ACGUTN" (probconsRNA/Defaults.h:28,
// wrapper.cpp:158-170).  U and T rows of the reference's tables are identical,
// so they share class 3; 4 = N; 5 = every other byte (1e-10 / 1e-5 emissions).
// Class 6 is internal padding: log-emission LOG_ZERO, used for columns past the
// end of a sequence so that padded DP cells can never win a log-sum.
enum { kClsN = 4, kClsOther = 5, kClsPad = 6, kNumCls = 7 };

inline uint8_t symbol_class(unsigned char c) {
  switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'U': case 'u': case 'T': case 't': return 3;
    case 'N': case 'n': return kClsN;
    default: return kClsOther;
  }
}

}  // namespace dafs
