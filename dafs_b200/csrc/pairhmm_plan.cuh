// The pair-HMM plan: shared by pairhmm.cu (which builds and runs it) and comm.cu (which
// hands its device-resident result to another rank).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <vector>

#include "../../include/dafs_cuda.h"

struct dafs_pairhmm_plan {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evk0 = nullptr, evk1 = nullptr;
  int nseq = 0, npairs = 0;
  float th = 0, xcut = 0;
  int C = 0, grid = 0;
  int64_t total_rows = 0;          // sum over pairs of (L1+1)
  std::vector<int64_t> row_off;    // host copy, [npairs+1]
  // device
  uint8_t* d_seq = nullptr; int64_t* d_seq_off = nullptr; int32_t* d_seq_len = nullptr;
  int2* d_pairs = nullptr; int32_t* d_order = nullptr;
  float* d_planes = nullptr; int64_t plane_stride = 0;
  float4* d_bnd = nullptr; int64_t bnd_stride = 0;
  int* d_queue = nullptr; unsigned long long* d_pool_count = nullptr;
  int64_t* d_row_off = nullptr; int32_t* d_row_ptr = nullptr;
  uint32_t* d_pool_col = nullptr; float* d_pool_val = nullptr; int64_t pool_cap = 0;
  int64_t* d_pair_pool_off = nullptr; int32_t* d_pair_nnz = nullptr; float* d_pair_total = nullptr;
  int64_t* d_nz_off = nullptr; uint32_t* d_col = nullptr; float* d_val = nullptr;
  // The pair list is run as up to kMaxChunks consecutive launches (whole rounds of the
  // resident warps each), so that fetch() can copy the CSR of a finished chunk to the
  // host while the next chunk computes.
  int nchunks = 1;
  int chunk_k0[5] = {0, 0, 0, 0, 0};
  cudaEvent_t ev_chunk[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t evck0[4] = {nullptr, nullptr, nullptr, nullptr}, evck1[4] = {nullptr, nullptr, nullptr, nullptr};
  int64_t* h_chunk_end = nullptr;   // pinned: nz_off at the end of each chunk
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_cp0 = nullptr, ev_cp1 = nullptr;   // around the result copies of a one-call run (copy rate, see g_copy_ratio)
  bool launched = false;
  bool consumed = false;            // the result arrays were given to a device batch
  int64_t nnz = -1;
  float last_ms = 0, last_kernel_ms = 0, last_copy_ms = 0;
};

