// Host side of (2): Boltzmann factors from the embedded BL* parameter blob, plan
// management, launch, CSR assembly and the C-ABI entry points declared in
// include/dafs_cuda.h.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "mccaskill_kernel.cuh"

namespace dafs {

// dafs_b200/data/blstar37.i32 (tools/extract_blstar.py), embedded at build time
static const int32_t kBlstar[] = {
#include "blstar37.inc"
};

// Reads the blob in the order of the reference's copy_* functions
// (src/boltzmann_param.c:5908-6026) and exponentiates at 37 C:
// kT = (37 + 273.15) * 1.98717 cal/mol, energies are in dcal/mol.
static void build_mc_tables(McTables& T) {
  std::memset(&T, 0, sizeof T);
  const int32_t* d = kBlstar;
  size_t p = 1;
  static int stack[8][8], mmH[8][5][5], mmI[8][5][5], d5[8][5], d3[8][5], i11[8][8][5][5], i21[8][8][5][5][5];
  static int i22[8][8][5][5][5][5];
  for (int i = 1; i <= 7; i++) for (int j = 1; j <= 7; j++) stack[i][j] = d[p++];
  for (int i = 1; i <= 7; i++) for (int a = 0; a < 5; a++) for (int b = 0; b < 5; b++) mmH[i][a][b] = d[p++];
  for (int i = 1; i <= 7; i++) for (int a = 0; a < 5; a++) for (int b = 0; b < 5; b++) mmI[i][a][b] = d[p++];
  for (int i = 0; i <= 7; i++) for (int a = 0; a < 5; a++) d5[i][a] = d[p++];
  for (int i = 0; i <= 7; i++) for (int a = 0; a < 5; a++) d3[i][a] = d[p++];
  for (int i = 1; i <= 7; i++) for (int j = 1; j <= 7; j++) for (int a = 0; a < 5; a++) for (int b = 0; b < 5; b++) i11[i][j][a][b] = d[p++];
  for (int i = 1; i <= 7; i++) for (int j = 1; j <= 7; j++) for (int a = 0; a < 5; a++) for (int b = 0; b < 5; b++) for (int c = 0; c < 5; c++) i21[i][j][a][b][c] = d[p++];
  for (int i = 1; i <= 7; i++) for (int j = 1; j <= 7; j++) for (int a = 1; a < 5; a++) for (int b = 1; b < 5; b++) for (int c = 1; c < 5; c++) for (int e = 1; e < 5; e++) i22[i][j][a][b][c][e] = d[p++];
  const double kT = (37.0 + 273.15) * 1.98717;
  auto bf = [&](int e) { return (float)std::exp(-(double)e * 10.0 / kT); };
  auto bfl = [&](int e) { return e >= 10000000 ? 0.0f : bf(e); };
  for (int i = 0; i < 31; i++) T.hairpin[i] = bfl(d[p++]);
  for (int i = 0; i < 31; i++) T.bulge[i] = bfl(d[p++]);
  for (int i = 0; i < 31; i++) T.internal[i] = bfl(d[p++]);
  const int ml_base = d[p++], ml_closing = d[p++], ml_intern = d[p++], termAU = d[p++];
  const int ninio = d[p++], max_ninio = d[p++];
  T.ml_base = bf(ml_base); T.ml_closing = bf(ml_closing); T.ml_intern = bf(ml_intern); T.termAU = bf(termAU);
  for (int k = 0; k <= MC_MAXLOOP; k++) T.ninio[k] = bf(std::min(k * ninio, max_ninio));
  T.lxc_over_kT = (float)(107.856 * 10.0 / kT);
  T.ntetra = std::min(d[p++], 32);
  for (int t = 0; t < T.ntetra; t++) {
    int code = 0;
    for (int c = 0; c < 6; c++) {
      const int ch = d[p++];
      const int nuc = ch == 'A' ? 1 : ch == 'C' ? 2 : ch == 'G' ? 3 : 4;
      code = code * 5 + nuc;
    }
    T.tetra_code[t] = code;
    T.tetra[t] = bf(d[p++]);
  }
  for (int i = 1; i <= 6; i++) {
    for (int j = 1; j <= 6; j++) {
      T.stack[i - 1][j - 1] = bf(stack[i][j]);
      for (int a = 0; a < 5; a++) for (int b = 0; b < 5; b++) {
        T.int11[i - 1][j - 1][a][b] = bf(i11[i][j][a][b]);
        for (int c = 0; c < 5; c++) T.int21[i - 1][j - 1][a][b][c] = bf(i21[i][j][a][b][c]);
      }
      for (int a = 1; a < 5; a++) for (int b = 1; b < 5; b++) for (int c = 1; c < 5; c++) for (int e = 1; e < 5; e++)
        T.int22[i - 1][j - 1][a - 1][b - 1][c - 1][e - 1] = bf(i22[i][j][a][b][c][e]);
    }
    for (int a = 0; a < 5; a++) {
      T.d5[i - 1][a] = bf(d5[i][a]); T.d3[i - 1][a] = bf(d3[i][a]);
      for (int b = 0; b < 5; b++) {
        T.mmH[i - 1][a][b] = bf(mmH[i][a][b]);
        T.mmI[i - 1][a][b] = bf(mmI[i][a][b]);
        // not in the BL* set: ViennaRNA's conversion rule for 1.8-format files
        T.mmM[i - 1][a][b] = bf(d5[i][a] + d3[i][b]);
        T.mmE[i - 1][a][b] = bf(d5[i][a] + d3[i][b]);
      }
    }
  }
}

static const McTables& mc_host_tables() {   // built once, thread-safe (function-local static)
  static const McTables* t = [] { McTables* x = new McTables(); build_mc_tables(*x); return x; }();
  return *t;
}

// ViennaRNA's nucleotide encoding: A C G U -> 1..4, T counts as U, anything else 0
static uint8_t mc_code(unsigned char c) {
  switch (c) {
    case 'A': case 'a': return 1;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 3;
    case 'U': case 'u': case 'T': case 't': return 4;
    default: return 0;
  }
}

__global__ void mc_scan_kernel(const int32_t* __restrict__ nnz, int n, int64_t* __restrict__ off) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    long long run = 0;
    for (int k = 0; k < n; ++k) { off[k] = run; run += nnz[k]; }
    off[n] = run;
  }
}

__global__ void mc_gather_kernel(int nseq, const int64_t* __restrict__ pool_off, const int32_t* __restrict__ nnz,
                                 const int64_t* __restrict__ nz_off, const uint32_t* __restrict__ pcol,
                                 const float* __restrict__ pval, uint32_t* __restrict__ col, float* __restrict__ val) {
  for (int s = blockIdx.x; s < nseq; s += gridDim.x) {
    const int64_t src = pool_off[s], dst = nz_off[s];
    const int n = nnz[s];
    for (int k = threadIdx.x; k < n; k += blockDim.x) { col[dst + k] = pcol[src + k]; val[dst + k] = pval[src + k]; }
  }
}

}  // namespace dafs

struct dafs_mccaskill_plan {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int nseq = 0, W = 0, grid = 0, smem = 0, cluster = 1;
  int gl = 8, threads = MC_THREADS;    // lanes per DP cell, threads per CTA (mccaskill_kernel<GL>)
  bool banded = false;                 // mccaskill_kernel<GL, true>
  float th = 0;
  int64_t total_rows = 0, pool_cap = 0, seq_bytes = 0;
  std::vector<int64_t> row_off;
  uint16_t* d_cons = nullptr;      // constraint classes | partners, each as long as d_seq (NULL: unconstrained)
  uint8_t* d_seq = nullptr; int64_t* d_seq_off = nullptr; int32_t* d_seq_len = nullptr; int32_t* d_order = nullptr;
  dafs::McTables* d_tab = nullptr;
  double* d_ws = nullptr; size_t ws_stride = 0;
  int64_t* d_row_off = nullptr; int32_t* d_row_ptr = nullptr;
  int64_t* d_pool_off = nullptr; uint32_t* d_pool_col = nullptr; float* d_pool_val = nullptr;
  int32_t* d_seq_nnz = nullptr; double* d_seq_lnz = nullptr;
  int64_t* d_nz_off = nullptr; uint32_t* d_col = nullptr; float* d_val = nullptr;
  bool launched = false;
  bool consumed = false;
  int64_t nnz = -1;
  float last_ms = 0;
};

namespace dafs {
static void free_mc_plan(dafs_mccaskill_plan* p) {
  dev_free(p->d_cons);
  dev_free(p->d_seq); dev_free(p->d_seq_off); dev_free(p->d_seq_len); dev_free(p->d_order); dev_free(p->d_tab);
  dev_free(p->d_ws); dev_free(p->d_row_off); dev_free(p->d_row_ptr); dev_free(p->d_pool_off);
  dev_free(p->d_pool_col); dev_free(p->d_pool_val); dev_free(p->d_seq_nnz); dev_free(p->d_seq_lnz);
  dev_free(p->d_nz_off); dev_free(p->d_col); dev_free(p->d_val);
  if (p->ev0) cudaEventDestroy(p->ev0);
  if (p->ev1) cudaEventDestroy(p->ev1);
  if (p->stream) cudaStreamDestroy(p->stream);
}
static const int kMcMaxLen = 6000;
// banded: the O(span) sums in bands of diagonals with a tiled far pass (mc_far_tile): for a sequence that one CTA
// of at least 256 threads owns; a cluster that shares a sequence keeps the cell-by-cell streams
static const void* mc_kernel_fn(int gl, bool banded) {
  if (banded) return gl == 8 ? (const void*)mccaskill_kernel<8, true> : gl == 16 ? (const void*)mccaskill_kernel<16, true> : (const void*)mccaskill_kernel<32, true>;
  return gl == 8 ? (const void*)mccaskill_kernel<8, false> : gl == 16 ? (const void*)mccaskill_kernel<16, false> : (const void*)mccaskill_kernel<32, false>;
}
}  // namespace dafs

using dafs::fail;

namespace dafs {
// Constraint string of one sequence (the reference's alphabet: '?' free, '.' stays unpaired, '(' ')'
// a given pair; anything else is treated as free) -> per-position class and partner (mc_allowed).
static int encode_constraint(const char* str, int n, uint16_t* cls, uint16_t* prt) {
  std::vector<int> open, open_loop;
  int next_loop = 1, cur = 0;
  for (int k = 0; k < n; ++k) {
    cls[k] = (uint16_t)cur; prt[k] = 0;
    const char c = str[k];
    if (c == '.') cls[k] = (uint16_t)(cur | 0x8000);
    else if (c == '(') {
      cls[k] = (uint16_t)(cur | 0x4000);
      open.push_back(k); open_loop.push_back(cur);
      if (next_loop >= 0x3fff) return -1;
      cur = next_loop++;
    } else if (c == ')') {
      if (open.empty()) return -1;
      const int a = open.back(); open.pop_back();
      cur = open_loop.back(); open_loop.pop_back();
      cls[k] = (uint16_t)(cur | 0x4000);
      prt[k] = (uint16_t)(a + 1); prt[a] = (uint16_t)(k + 1);
    }
  }
  return open.empty() ? 0 : -1;
}
}  // namespace dafs

extern "C" int dafs_cuda_mccaskill_plan_create(int nseq, const char* const* seq, const int* len, float th,
                                               dafs_mccaskill_plan** out) {
  return dafs_cuda_mccaskill_plan_create_constrained(nseq, seq, len, nullptr, th, out);
}

extern "C" int dafs_cuda_mccaskill_plan_create_constrained(int nseq, const char* const* seq, const int* len,
                                                           const char* const* str, float th,
                                                           dafs_mccaskill_plan** out) {
  using namespace dafs;
  if (!out) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  *out = nullptr;
  if (nseq < 0 || (nseq > 0 && (!seq || !len))) return fail(DAFS_CUDA_EINVAL, "bad nseq / NULL input array");
  if (std::isnan(th)) return fail(DAFS_CUDA_EINVAL, "threshold is NaN");
  int maxn = 1;
  for (int i = 0; i < nseq; ++i) {
    if (len[i] < 1 || !seq[i]) return fail(DAFS_CUDA_EINVAL, "sequence %d is empty", i);
    if (len[i] > kMcMaxLen) return fail(DAFS_CUDA_ENOTSUP, "sequence %d has %d nt; this build folds up to %d", i, len[i], kMcMaxLen);
    maxn = std::max(maxn, len[i]);
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  }
  dafs_mccaskill_plan* p = new dafs_mccaskill_plan();
  auto bail = [&](int rc) { free_mc_plan(p); delete p; return rc; };
#define PLAN_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return bail(fail(e_ == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM : DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_))); } while (0)
  PLAN_TRY(cudaGetDevice(&p->device));
  p->nseq = nseq; p->th = th; p->W = maxn + 2;
  PLAN_TRY(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
  PLAN_TRY(cudaEventCreate(&p->ev0));
  PLAN_TRY(cudaEventCreate(&p->ev1));
  if (nseq == 0) { *out = p; return DAFS_CUDA_OK; }

  std::vector<int64_t> seq_off(nseq + 1, 0), pool_off(nseq + 1, 0);
  p->row_off.assign(nseq + 1, 0);
  // a row holds at most floor(1/th) entries above th (its probabilities sum to <= 1)
  const int64_t per_row = (th > 0.0f) ? (int64_t)std::floor(1.0 / (double)th) + 2 : (int64_t)maxn;
  for (int i = 0; i < nseq; ++i) {
    seq_off[i + 1] = seq_off[i] + len[i];
    p->row_off[i + 1] = p->row_off[i] + len[i] + 1;
    int64_t cap = 0;
    for (int r = 0; r < len[i]; ++r) cap += std::min<int64_t>(per_row, len[i] - 1 - r);
    pool_off[i + 1] = pool_off[i] + cap + 8;
  }
  p->total_rows = p->row_off[nseq]; p->pool_cap = pool_off[nseq];
  std::vector<uint8_t> enc(seq_off[nseq]);
  p->seq_bytes = seq_off[nseq];
  for (int i = 0; i < nseq; ++i)
    for (int k = 0; k < len[i]; ++k) enc[seq_off[i] + k] = mc_code((unsigned char)seq[i][k]);
  std::vector<int32_t> order(nseq);
  for (int i = 0; i < nseq; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return len[a] > len[b]; });

  dafs::DevInfo prop;
  PLAN_TRY(dafs::dev_info(p->device, &prop));
  const size_t tab_bytes = (sizeof(McTables) + 15) & ~size_t(15);
  p->smem = (int)(tab_bytes + sizeof(McCoef) + sizeof(double) * MC_FAR_STAGE + 2 * sizeof(double) * (p->W + 2) + 2 * sizeof(uint16_t) * p->W + p->W + 16 + 2 * sizeof(uint16_t) * (p->W + 2));
  if ((size_t)p->smem > prop.smem_optin)
    return bail(fail(DAFS_CUDA_ENOTSUP, "sequence of %d nt needs %d bytes of shared memory", maxn, p->smem));
  // Lanes per DP cell: the sums of a cell are as long as its span, so short sequences (whose
  // diagonals hold fewer cells than the CTA has lanes) get 8-lane groups and long ones whole warps.
  p->gl = maxn <= 480 ? 8 : maxn <= 1100 ? 16 : 32;
  if (const char* e = getenv("DAFS_MCCASKILL_LANES")) { const int v = atoi(e); if (v == 8 || v == 16 || v == 32) p->gl = v; }
  // Throughput form (at least one sequence per SM): two CTAs of 384 threads per SM, each with a
  // sequence of its own, so that one sequence's per-diagonal barrier and serial cell tails are
  // covered by the other's work.  Latency form (fewer sequences than SMs): one CTA of 768 threads,
  // in a thread-block cluster per sequence.
  p->threads = (nseq >= prop.sms && 2 * (size_t)p->smem + 2048 <= prop.smem_optin) ? MC_THREADS / 2 : MC_THREADS;
  if (const char* e = getenv("DAFS_MCCASKILL_THREADS")) { const int v = atoi(e); if (v >= 64 && v <= MC_THREADS && v % 32 == 0) p->threads = v; }
  const void* fn = mc_kernel_fn(p->gl, true);        // (both forms: same shared memory, same register bound)
  PLAN_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem));
  int blocks = 0;
  PLAN_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, fn, p->threads, p->smem));
  if (blocks < 1) return bail(fail(DAFS_CUDA_ECUDA, "mccaskill_kernel does not fit on an SM"));
  p->ws_stride = (size_t)MC_NARRAYS * p->W * p->W + (size_t)MC_NROWS * p->W;
  size_t free_b = 0, total_b = 0;
  PLAN_TRY(cudaMemGetInfo(&free_b, &total_b));
  const size_t per_cta = p->ws_stride * sizeof(double);
  const int by_mem = (int)std::max<size_t>(1, (free_b * 6 / 10) / per_cta);
  // CTAs per sequence: a lone CTA when there is at least one sequence per SM slot,
  // otherwise the largest portable cluster (2, 4, 8) that still gives every
  // sequence a team; DAFS_MCCASKILL_CLUSTER overrides (tuning knob).
  const int slots = prop.sms * blocks;
  int nc = 1;
  while (nc < 8 && (int64_t)nseq * nc * 2 <= slots) nc *= 2;
  // Very few sequences (a C5 shard of an 8-GPU run holds two): a non-portable cluster of 16 CTAs per sequence, when
  // the device can co-schedule one per sequence (a cluster lives inside one GPC)
  bool allow16 = false;
  if (nc == 8 && (int64_t)nseq * 16 * 2 <= slots) {
    const void* fs = mc_kernel_fn(p->gl, false);
    if (cudaFuncSetAttribute(fs, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
        cudaFuncSetAttribute(fs, cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem) == cudaSuccess) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(16 * nseq); cfg.blockDim = dim3(p->threads); cfg.dynamicSmemBytes = p->smem;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 16; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      int ncl = 0;
      if (cudaOccupancyMaxActiveClusters(&ncl, fs, &cfg) == cudaSuccess && ncl >= nseq) { nc = 16; allow16 = true; }
    }
    (void)cudaGetLastError();
  }
  if (const char* e = getenv("DAFS_MCCASKILL_CLUSTER")) {
    const int v = atoi(e);
    if (v == 1 || v == 2 || v == 4 || v == 8) nc = v;
    else if (v == 16 && allow16) nc = 16;
  }
  p->cluster = nc;
  p->banded = nc == 1 && p->threads >= 256;
  const int teams = std::max(1, std::min(std::min(nseq, slots / nc), by_mem));
  p->grid = teams * nc;

  PLAN_TRY(dev_malloc(&p->d_seq, enc.size()));
  PLAN_TRY(dev_malloc(&p->d_seq_off, sizeof(int64_t) * (nseq + 1)));
  PLAN_TRY(dev_malloc(&p->d_seq_len, sizeof(int32_t) * nseq));
  PLAN_TRY(dev_malloc(&p->d_order, sizeof(int32_t) * nseq));
  PLAN_TRY(dev_malloc(&p->d_tab, sizeof(McTables)));
  PLAN_TRY(dev_malloc(&p->d_ws, per_cta * (p->grid / p->cluster)));
  PLAN_TRY(dev_malloc(&p->d_row_off, sizeof(int64_t) * (nseq + 1)));
  PLAN_TRY(dev_malloc(&p->d_row_ptr, sizeof(int32_t) * p->total_rows));
  PLAN_TRY(dev_malloc(&p->d_pool_off, sizeof(int64_t) * (nseq + 1)));
  PLAN_TRY(dev_malloc(&p->d_pool_col, sizeof(uint32_t) * p->pool_cap));
  PLAN_TRY(dev_malloc(&p->d_pool_val, sizeof(float) * p->pool_cap));
  PLAN_TRY(dev_malloc(&p->d_seq_nnz, sizeof(int32_t) * nseq));
  PLAN_TRY(dev_malloc(&p->d_seq_lnz, sizeof(double) * nseq));
  PLAN_TRY(dev_malloc(&p->d_nz_off, sizeof(int64_t) * (nseq + 1)));
  PLAN_TRY(dev_malloc(&p->d_col, sizeof(uint32_t) * p->pool_cap));
  PLAN_TRY(dev_malloc(&p->d_val, sizeof(float) * p->pool_cap));
  if (str) {
    const size_t tot = enc.size();
    std::vector<uint16_t> cons(2 * tot, 0);
    for (int i = 0; i < nseq; ++i) {
      if (!str[i]) { continue; }       // this sequence is unconstrained: class 0 (one loop), no partners
      if (encode_constraint(str[i], len[i], cons.data() + seq_off[i], cons.data() + tot + seq_off[i]))
        return bail(fail(DAFS_CUDA_EINVAL, "sequence %d: unbalanced brackets in the constraint", i));
    }
    PLAN_TRY(dev_malloc(&p->d_cons, sizeof(uint16_t) * 2 * std::max<size_t>(tot, 1)));
    PLAN_TRY(cudaMemcpyAsync(p->d_cons, cons.data(), sizeof(uint16_t) * 2 * tot, cudaMemcpyHostToDevice, p->stream));
    PLAN_TRY(cudaStreamSynchronize(p->stream));
  }
  PLAN_TRY(cudaMemcpyAsync(p->d_seq, enc.data(), enc.size(), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_seq_off, seq_off.data(), sizeof(int64_t) * (nseq + 1), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_seq_len, len, sizeof(int32_t) * nseq, cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_order, order.data(), sizeof(int32_t) * nseq, cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_tab, &mc_host_tables(), sizeof(McTables), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_row_off, p->row_off.data(), sizeof(int64_t) * (nseq + 1), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_pool_off, pool_off.data(), sizeof(int64_t) * (nseq + 1), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaStreamSynchronize(p->stream));
#undef PLAN_TRY
  *out = p;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_mccaskill_plan_run(dafs_mccaskill_plan* p) {
  using namespace dafs;
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result arrays were handed over (dafs_cuda_mccaskill_plan_fetch_device): create a new plan");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  p->launched = true; p->nnz = -1;
  if (p->nseq == 0) { p->nnz = 0; return DAFS_CUDA_OK; }
  McArgs a;
  a.seq = p->d_seq; a.seq_off = p->d_seq_off; a.seq_len = p->d_seq_len; a.order = p->d_order; a.nseq = p->nseq;
  a.cons_cls = p->d_cons; a.cons_prt = p->d_cons ? p->d_cons + p->seq_bytes : nullptr;
  a.W = p->W; a.th = p->th; a.tab = p->d_tab; a.ws = p->d_ws; a.ws_stride = p->ws_stride; a.cluster = p->cluster;
  a.row_off = p->d_row_off; a.row_ptr = p->d_row_ptr; a.pool_off = p->d_pool_off; a.pool_col = p->d_pool_col;
  a.pool_val = p->d_pool_val; a.seq_nnz = p->d_seq_nnz; a.seq_lnz = p->d_seq_lnz;
  DAFS_CUDA_TRY(cudaEventRecord(p->ev0, p->stream));
  const void* fn = mc_kernel_fn(p->gl, p->banded);
  DAFS_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem));
  if (p->cluster > 8) DAFS_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p->grid); cfg.blockDim = dim3(p->threads);
    cfg.dynamicSmemBytes = p->smem; cfg.stream = p->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = p->cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    void* kargs[] = {(void*)&a};
    DAFS_CUDA_TRY(cudaLaunchKernelExC(&cfg, fn, kargs));
  }
  mc_scan_kernel<<<1, 32, 0, p->stream>>>(p->d_seq_nnz, p->nseq, p->d_nz_off);
  DAFS_CUDA_TRY(cudaGetLastError());
  mc_gather_kernel<<<std::min(p->nseq, 1024), 256, 0, p->stream>>>(p->nseq, p->d_pool_off, p->d_seq_nnz, p->d_nz_off,
                                                                 p->d_pool_col, p->d_pool_val, p->d_col, p->d_val);
  DAFS_CUDA_TRY(cudaGetLastError());
  DAFS_CUDA_TRY(cudaEventRecord(p->ev1, p->stream));
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_mccaskill_plan_sync(dafs_mccaskill_plan* p, float* device_ms, int64_t* nnz) {
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  DAFS_CUDA_TRY(cudaStreamSynchronize(p->stream));
  if (p->nseq > 0 && p->nnz < 0) {
    int64_t total = 0;
    DAFS_CUDA_TRY(cudaMemcpy(&total, p->d_nz_off + p->nseq, sizeof total, cudaMemcpyDeviceToHost));
    p->nnz = total;
    DAFS_CUDA_TRY(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
  }
  if (device_ms) *device_ms = p->last_ms;
  if (nnz) *nnz = p->nnz;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_mccaskill_plan_fetch(dafs_mccaskill_plan* p, dafs_csr_batch* out, double* lnz) {
  if (!p || !out) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  *out = dafs_csr_batch{};
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was handed over to a device batch");
  const int rc = dafs_cuda_mccaskill_plan_sync(p, nullptr, nullptr);
  if (rc) return rc;
  const int64_t ns = p->nseq, nnz = p->nnz, rows = p->total_rows;
  const size_t b_off = sizeof(int64_t) * (ns + 1);
  const size_t b_row_ptr = sizeof(int32_t) * rows, b_col = sizeof(uint32_t) * nnz, b_val = sizeof(float) * nnz;
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t total = 2 * up(b_off) + up(b_row_ptr) + up(b_col) + up(b_val);
  char* base = (char*)dafs::pinned_alloc(total);
  if (!base) return fail(DAFS_CUDA_ENOMEM, "pinned allocation of %zu bytes failed", total);
  int64_t* h_row_off = (int64_t*)base;
  int64_t* h_nz_off = (int64_t*)(base + up(b_off));
  int32_t* h_row_ptr = (int32_t*)((char*)h_nz_off + up(b_off));
  uint32_t* h_col = (uint32_t*)((char*)h_row_ptr + up(b_row_ptr));
  float* h_val = (float*)((char*)h_col + up(b_col));
  if (ns > 0) std::memcpy(h_row_off, p->row_off.data(), b_off); else h_row_off[0] = 0;
  cudaError_t e = cudaSuccess;
  if (ns > 0) {
    e = cudaMemcpyAsync(h_nz_off, p->d_nz_off, b_off, cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_row_ptr, p->d_row_ptr, b_row_ptr, cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess && nnz) e = cudaMemcpyAsync(h_col, p->d_col, b_col, cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess && nnz) e = cudaMemcpyAsync(h_val, p->d_val, b_val, cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess && lnz) e = cudaMemcpyAsync(lnz, p->d_seq_lnz, sizeof(double) * ns, cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
  } else {
    h_nz_off[0] = 0;
  }
  if (e != cudaSuccess) { dafs::pinned_free(base); return fail(DAFS_CUDA_ECUDA, "result copy failed: %s", cudaGetErrorString(e)); }
  out->n_units = ns; out->nnz = nnz; out->row_off = h_row_off; out->row_ptr = h_row_ptr; out->nz_off = h_nz_off;
  out->col = h_col; out->val = h_val; out->owner_ = base;
  return DAFS_CUDA_OK;
}

// The result stays in HBM (see dafs_cuda_pairhmm_plan_fetch_device): the plan gives up its row_ptr /
// col / val arrays and cannot be run again.
extern "C" int dafs_cuda_mccaskill_plan_fetch_device(dafs_mccaskill_plan* p, dafs_csr_batch* out) {
  if (!p || !out) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  *out = dafs_csr_batch{};
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was already handed over");
  const int rc = dafs_cuda_mccaskill_plan_sync(p, nullptr, nullptr);
  if (rc) return rc;
  if (p->nseq == 0) return fail(DAFS_CUDA_EINVAL, "empty plan");
  const int rc2 = dafs::make_device_batch(p->nseq, p->nnz, p->row_off.data(), p->d_nz_off, p->d_row_ptr, p->d_col, p->d_val, p->stream, out);
  if (rc2) return rc2;
  p->d_row_ptr = nullptr; p->d_col = nullptr; p->d_val = nullptr;
  p->consumed = true;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_mccaskill_plan_info(dafs_mccaskill_plan* p, dafs_mccaskill_info* info) {
  if (!p || !info) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  info->grid = p->grid; info->block = p->threads; info->smem_bytes = p->smem; info->row_stride = p->W;
  info->workspace_bytes = (int64_t)(p->ws_stride * sizeof(double)) * (p->grid / p->cluster);
  info->table_bytes = (int)sizeof(dafs::McTables);
  info->last_ms = p->last_ms;
  return DAFS_CUDA_OK;
}

extern "C" void dafs_cuda_mccaskill_plan_destroy(dafs_mccaskill_plan* p) {
  if (!p) return;
  cudaSetDevice(p->device);
  if (p->stream) cudaStreamSynchronize(p->stream);
  dafs::free_mc_plan(p);
  delete p;
}

// RNAfold::calculate(seq, str, bp) for a batch (src/fold.cpp:118-166): one device, one call.
extern "C" int dafs_cuda_mccaskill_constrained(int nseq, const char* const* seq, const int* len,
                                               const char* const* str, float th, dafs_csr_batch* out) {
  if (!out) return fail(DAFS_CUDA_EINVAL, "out is NULL");
  *out = dafs_csr_batch{};
  dafs_mccaskill_plan* p = nullptr;
  int rc = dafs_cuda_mccaskill_plan_create_constrained(nseq, seq, len, str, th, &p);
  if (rc) return rc;
  rc = dafs_cuda_mccaskill_plan_run(p);
  if (!rc) rc = dafs_cuda_mccaskill_plan_fetch(p, out, nullptr);
  dafs_cuda_mccaskill_plan_destroy(p);
  return rc;
}

// All sequences (src/fold.cpp:60-68).  ngpu > 1: sequences dealt round-robin by
// descending length over the devices of this process, merged on the host.
extern "C" int dafs_cuda_mccaskill_all(int nseq, const char* const* seq, const int* len, float th, int ngpu,
                                       dafs_csr_batch* out) {
  if (!out) return fail(DAFS_CUDA_EINVAL, "out is NULL");
  *out = dafs_csr_batch{};
  const int ndev = dafs_cuda_device_count();
  if (ndev <= 0) return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  if (ngpu > ndev) return fail(DAFS_CUDA_ENODEV, "requested %d GPUs, %d visible", ngpu, ndev);
  if (ngpu <= 1 || nseq < 2 * ngpu) {
    dafs_mccaskill_plan* p = nullptr;
    int rc = dafs_cuda_mccaskill_plan_create(nseq, seq, len, th, &p);
    if (rc) return rc;
    rc = dafs_cuda_mccaskill_plan_run(p);
    if (!rc) rc = dafs_cuda_mccaskill_plan_fetch(p, out, nullptr);
    dafs_cuda_mccaskill_plan_destroy(p);
    return rc;
  }
  std::vector<int> idx(nseq);
  for (int i = 0; i < nseq; ++i) idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return len[a] > len[b]; });
  std::vector<std::vector<const char*>> sseq(ngpu);
  std::vector<std::vector<int>> slen(ngpu), sidx(ngpu);
  for (int k = 0; k < nseq; ++k) { const int d = k % ngpu; sseq[d].push_back(seq[idx[k]]); slen[d].push_back(len[idx[k]]); sidx[d].push_back(idx[k]); }
  std::vector<dafs_csr_batch> part(ngpu);
  std::vector<int> rc(ngpu, 0);
  std::vector<std::string> err(ngpu);
  std::vector<std::thread> workers;
  for (int d = 0; d < ngpu; ++d)
    workers.emplace_back([&, d]() {
      if (cudaSetDevice(d) != cudaSuccess) { rc[d] = DAFS_CUDA_ECUDA; err[d] = "cudaSetDevice failed"; return; }
      rc[d] = dafs_cuda_mccaskill_all((int)sseq[d].size(), sseq[d].data(), slen[d].data(), th, 1, &part[d]);
      if (rc[d]) err[d] = dafs::last_error();
    });
  for (auto& w : workers) w.join();
  for (int d = 0; d < ngpu; ++d)
    if (rc[d]) { for (auto& b : part) dafs_cuda_free_batch(&b); return fail(rc[d], "device %d: %s", d, err[d].c_str()); }
  std::vector<std::pair<int, int>> where(nseq);      // sequence -> (device, unit)
  for (int d = 0; d < ngpu; ++d) for (size_t u = 0; u < sidx[d].size(); ++u) where[sidx[d][u]] = {d, (int)u};
  int64_t nnz = 0, rows = 0;
  for (int i = 0; i < nseq; ++i) rows += len[i] + 1;
  for (auto& b : part) nnz += b.nnz;
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t b_off = sizeof(int64_t) * (nseq + 1);
  const size_t total = 2 * up(b_off) + up(sizeof(int32_t) * rows) + up(sizeof(uint32_t) * nnz) + up(sizeof(float) * nnz);
  char* base = (char*)dafs::pinned_alloc(total);
  if (!base) { for (auto& b : part) dafs_cuda_free_batch(&b); return fail(DAFS_CUDA_ENOMEM, "pinned allocation failed"); }
  int64_t* row_off = (int64_t*)base;
  int64_t* nz_off = (int64_t*)(base + up(b_off));
  int32_t* row_ptr = (int32_t*)((char*)nz_off + up(b_off));
  uint32_t* col = (uint32_t*)((char*)row_ptr + up(sizeof(int32_t) * rows));
  float* val = (float*)((char*)col + up(sizeof(uint32_t) * nnz));
  row_off[0] = 0; nz_off[0] = 0;
  for (int i = 0; i < nseq; ++i) {
    const dafs_csr_batch& b = part[where[i].first];
    const int64_t u = where[i].second;
    const int64_t nr = b.row_off[u + 1] - b.row_off[u], nz = b.nz_off[u + 1] - b.nz_off[u];
    std::memcpy(row_ptr + row_off[i], b.row_ptr + b.row_off[u], sizeof(int32_t) * nr);
    std::memcpy(col + nz_off[i], b.col + b.nz_off[u], sizeof(uint32_t) * nz);
    std::memcpy(val + nz_off[i], b.val + b.nz_off[u], sizeof(float) * nz);
    row_off[i + 1] = row_off[i] + nr; nz_off[i + 1] = nz_off[i] + nz;
  }
  for (auto& b : part) dafs_cuda_free_batch(&b);
  out->n_units = nseq; out->nnz = nnz; out->row_off = row_off; out->row_ptr = row_ptr; out->nz_off = nz_off;
  out->col = col; out->val = val; out->owner_ = base;
  return DAFS_CUDA_OK;
}
