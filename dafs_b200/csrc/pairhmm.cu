// Host side of (1): plan management, launch, CSR assembly and the C-ABI entry
// points declared in include/dafs_cuda.h.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <numeric>
#include <mutex>
#include <thread>
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "pairhmm_plan.cuh"
#include "pairhmm_post.cuh"

namespace dafs {

// ---- model tables: ProbabilisticModel ctor (ProbabilisticModel.h:55-88) over
// the defaults of probconsRNA/Defaults.h:19-39 --------------------------------
static const float kInit[3] = {0.9588437676f, 0.0205782652f, 0.0205782652f};
static const float kGapOpen = 0.0190259293f, kGapExtend = 0.3269913495f;
static const float kEmitSingle[5] = {0.2270790040f, 0.2422080040f, 0.2839320004f, 0.2464679927f,
                                     0.0003124650f};   // A C G U/T N
static const float kEmitPairs[5][5] = {
    {0.1487240046f, 0.0184142999f, 0.0361397006f, 0.0238473993f, 0.0000375308f},
    {0.0184142999f, 0.1583919972f, 0.0275536999f, 0.0389291011f, 0.0000815823f},
    {0.0361397006f, 0.0275536999f, 0.1979320049f, 0.0244289003f, 0.0000824765f},
    {0.0238473993f, 0.0389291011f, 0.0244289003f, 0.1557479948f, 0.0000743985f},
    {0.0000375308f, 0.0000815823f, 0.0000824765f, 0.0000743985f, 0.0000263252f}};

static void build_tables(PairHmmTables& t) {
  float tm[3][3] = {{0.f}};
  tm[0][0] = 1;
  tm[0][1] = kGapOpen; tm[0][2] = kGapOpen;
  tm[0][0] -= (kGapOpen + kGapOpen);
  tm[1][1] = kGapExtend; tm[2][2] = kGapExtend;
  tm[1][0] = 1 - kGapExtend; tm[2][0] = 1 - kGapExtend;
  for (int i = 0; i < 3; ++i) {
    t.init[i] = logf(kInit[i]);
    for (int j = 0; j < 3; ++j) t.trans[i * 3 + j] = logf(tm[i][j]);   // log 0 = -inf, never read
  }
  for (int a = 0; a < 8; ++a) {
    t.ins[a] = a < 5 ? logf(kEmitSingle[a]) : a == kClsOther ? logf((float)1e-5) : DAFS_LOG_ZERO;
    for (int b = 0; b < 8; ++b) {
      float v;
      if (a < 5 && b < 5) v = logf(kEmitPairs[a][b]);
      else if (a <= kClsOther && b <= kClsOther) v = logf((float)1e-10);
      else v = DAFS_LOG_ZERO;
      t.match[a * 8 + b] = v;
    }
  }
}

// built once, by whichever thread comes first (function-local static: the ngpu > 1 entry
// points create plans from one host thread per device)
static const PairHmmTables& host_tables() {
  static const PairHmmTables t = [] { PairHmmTables x; build_tables(x); return x; }();
  return t;
}

// keep-threshold prefilter: x <= xcut implies EXP(x) <= th (EXP's pieces are
// within ~1.1e-4 of exp on [-16,0]), so such cells are never kept.
static float compute_xcut(float th) {
  if (th < 0.f) return -INFINITY;
  if (th < 5.2e-4f) return -16.0f;
  return logf(th) - 0.5f;
}

// ---- small helper kernels ---------------------------------------------------
// Exclusive scan of per-pair nnz -> nz_off[npairs+1]; one block.
// has_base: the range continues a scan, off[0] already holds its start (chunked runs).
__global__ void scan_nnz_kernel(const int32_t* __restrict__ nnz, int n, int64_t* __restrict__ off, int has_base) {
  __shared__ long long part[1024];
  const int t = threadIdx.x, T = blockDim.x;
  const long long base = has_base ? off[0] : 0;
  __syncthreads();
  const int per = (n + T - 1) / T;
  const int b = min(n, t * per), e = min(n, b + per);
  long long s = 0;
  for (int k = b; k < e; ++k) s += nnz[k];
  part[t] = s;
  __syncthreads();
  if (t == 0) {
    long long run = base;
    for (int k = 0; k < T; ++k) { const long long v = part[k]; part[k] = run; run += v; }
    off[n] = run;
  }
  __syncthreads();
  long long run = part[t];
  for (int k = b; k < e; ++k) { off[k] = run; run += nnz[k]; }
}

// Pool (completion order) -> pair order.  One warp per pair.  When the pool was
// too small (the host sees pool_count > cap, grows it and runs again) the pairs
// that did not fit are skipped: neither side may be touched past cap.
__global__ void gather_csr_kernel(int npairs, const int64_t* __restrict__ pool_off,
                                  const int32_t* __restrict__ nnz, const int64_t* __restrict__ nz_off,
                                  const uint32_t* __restrict__ pcol, const float* __restrict__ pval,
                                  uint32_t* __restrict__ col, float* __restrict__ val, int64_t cap) {
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  for (int p = blockIdx.x * wpb + (threadIdx.x >> 5); p < npairs; p += gridDim.x * wpb) {
    const int64_t src = pool_off[p], dst = nz_off[p];
    const int n = nnz[p];
    if (src + n > cap || dst + n > cap) continue;
    for (int k = lane; k < n; k += 32) { col[dst + k] = pcol[src + k]; val[dst + k] = pval[src + k]; }
  }
}

}  // namespace dafs

// ---- plan (struct dafs_pairhmm_plan: pairhmm_plan.cuh) -------------------------
namespace dafs {

static const int kWarps = 4;     // warps per CTA
#ifndef DAFS_PAIRHMM_MINB
#define DAFS_PAIRHMM_MINB 4
#endif
static const int kMinBlocks = DAFS_PAIRHMM_MINB; // CTAs per SM the kernel is compiled for (4 -> <=128 regs)

static void free_plan_buffers(dafs_pairhmm_plan* p) {
  dev_free(p->d_seq); dev_free(p->d_seq_off); dev_free(p->d_seq_len); dev_free(p->d_pairs);
  dev_free(p->d_order); dev_free(p->d_planes); dev_free(p->d_bnd); dev_free(p->d_queue);
  dev_free(p->d_pool_count); dev_free(p->d_row_off); dev_free(p->d_row_ptr);
  dev_free(p->d_pool_col); dev_free(p->d_pool_val); dev_free(p->d_pair_pool_off);
  dev_free(p->d_pair_nnz); dev_free(p->d_pair_total); dev_free(p->d_nz_off);
  dev_free(p->d_col); dev_free(p->d_val);
  if (p->ev0) cudaEventDestroy(p->ev0);
  if (p->ev1) cudaEventDestroy(p->ev1);
  if (p->evk0) cudaEventDestroy(p->evk0);
  if (p->evk1) cudaEventDestroy(p->evk1);
  if (p->ev_cp0) cudaEventDestroy(p->ev_cp0);
  if (p->ev_cp1) cudaEventDestroy(p->ev_cp1);
  for (int c = 0; c < 4; ++c) {
    if (p->ev_chunk[c]) cudaEventDestroy(p->ev_chunk[c]);
    if (p->evck0[c]) cudaEventDestroy(p->evck0[c]);
    if (p->evck1[c]) cudaEventDestroy(p->evck1[c]);
  }
  if (p->h_chunk_end) pinned_free(p->h_chunk_end);
  if (p->copy_stream) cudaStreamDestroy(p->copy_stream);
  if (p->stream) cudaStreamDestroy(p->stream);
}

static int alloc_pool(dafs_pairhmm_plan* p, int64_t cap) {
  dev_free(p->d_pool_col); dev_free(p->d_pool_val); dev_free(p->d_col); dev_free(p->d_val);
  p->d_pool_col = nullptr; p->d_pool_val = nullptr; p->d_col = nullptr; p->d_val = nullptr;
  cap = std::max<int64_t>(cap, 32);
  DAFS_CUDA_TRY(dev_malloc(&p->d_pool_col, cap * sizeof(uint32_t)));
  DAFS_CUDA_TRY(dev_malloc(&p->d_pool_val, cap * sizeof(float)));
  DAFS_CUDA_TRY(dev_malloc(&p->d_col, cap * sizeof(uint32_t)));
  DAFS_CUDA_TRY(dev_malloc(&p->d_val, cap * sizeof(float)));
  p->pool_cap = cap;
  return DAFS_CUDA_OK;
}

// Columns per lane: minimise (padded DP width) x (measured cost per padded cell of
// that instantiation).  The relative costs are from the C4 sweep on B200
// (tools/sweep_pairhmm.sh, profiles/r01_pairhmm_cols_sweep.txt): wide lanes
// amortise the per-step shuffles, narrow ones keep the unrolled row body inside
// the instruction cache.
static const int kColsPerLane[] = {2, 3, 4, 5, 6, 8, 10, 12};
static const float kColsCost[] = {1.35f, 1.15f, 1.10f, 1.00f, 0.975f, 1.10f, 1.25f, 1.03f};   // time per padded column, B200, round 2
static int choose_cols(int max_l2) {
  if (const char* e = getenv("DAFS_PAIRHMM_COLS")) {          // tuning knob
    const int want = atoi(e);
    for (int c : kColsPerLane) if (c == want) return c;
  }
  int best = 5; double best_w = 1e300;
  for (size_t k = 0; k < sizeof(kColsPerLane) / sizeof(int); ++k) {
    const int sw = 32 * kColsPerLane[k];
    const double w = (double)((max_l2 + 1 + sw - 1) / sw) * sw * kColsCost[k];
    if (w < best_w) { best_w = w; best = kColsPerLane[k]; }
  }
  return best;
}

template <int C>
static cudaError_t launch_pairhmm(const PairHmmArgs& a, int grid, cudaStream_t st) {
  pairhmm_kernel<C, kWarps, kMinBlocks><<<grid, kWarps * 32, 0, st>>>(a);
  return cudaGetLastError();
}

static int launch_all(dafs_pairhmm_plan* p) {
  PairHmmArgs a;
  a.seq = p->d_seq; a.seq_off = p->d_seq_off; a.seq_len = p->d_seq_len;
  a.pairs = p->d_pairs; a.order = p->d_order; a.npairs = p->npairs;
  a.th = p->th; a.xcut = p->xcut;
  a.planes = p->d_planes; a.plane_stride = p->plane_stride;
  a.bnd = p->d_bnd; a.bnd_stride = p->bnd_stride;
  a.queue = p->d_queue;
  a.row_off = p->d_row_off; a.row_ptr = p->d_row_ptr;
  a.pool_count = p->d_pool_count; a.pool_col = p->d_pool_col; a.pool_val = p->d_pool_val;
  a.pool_cap = p->pool_cap;
  a.pair_pool_off = p->d_pair_pool_off; a.pair_nnz = p->d_pair_nnz; a.pair_total = p->d_pair_total;
  DAFS_CUDA_TRY(cudaMemsetAsync(p->d_pool_count, 0, sizeof(unsigned long long), p->stream));
  DAFS_CUDA_TRY(cudaEventRecord(p->ev0, p->stream));
  for (int c = 0; c < p->nchunks; ++c) {
    const int k0 = p->chunk_k0[c], kn = p->chunk_k0[c + 1] - k0;
    a.order = p->d_order + k0; a.npairs = kn;
    DAFS_CUDA_TRY(cudaMemsetAsync(p->d_queue, 0, sizeof(int), p->stream));
    DAFS_CUDA_TRY(cudaEventRecord(p->evck0[c], p->stream));
    const int grid = (int)std::min<int64_t>(p->grid, (kn + kWarps - 1) / kWarps);
    cudaError_t e;
    switch (p->C) {
      case 2: e = launch_pairhmm<2>(a, grid, p->stream); break;
      case 3: e = launch_pairhmm<3>(a, grid, p->stream); break;
      case 4: e = launch_pairhmm<4>(a, grid, p->stream); break;
      case 5: e = launch_pairhmm<5>(a, grid, p->stream); break;
      case 6: e = launch_pairhmm<6>(a, grid, p->stream); break;
      case 8: e = launch_pairhmm<8>(a, grid, p->stream); break;
      case 10: e = launch_pairhmm<10>(a, grid, p->stream); break;
      default: e = launch_pairhmm<12>(a, grid, p->stream); break;
    }
    DAFS_CUDA_TRY(e);
    DAFS_CUDA_TRY(cudaEventRecord(p->evck1[c], p->stream));
    scan_nnz_kernel<<<1, 1024, 0, p->stream>>>(p->d_pair_nnz + k0, kn, p->d_nz_off + k0, c > 0);
    DAFS_CUDA_TRY(cudaGetLastError());
    const int gblocks = std::min(kn / 8 + 1, 148 * 8);
    gather_csr_kernel<<<gblocks, 256, 0, p->stream>>>(kn, p->d_pair_pool_off + k0, p->d_pair_nnz + k0,
                                                      p->d_nz_off + k0, p->d_pool_col, p->d_pool_val,
                                                      p->d_col, p->d_val, p->pool_cap);
    DAFS_CUDA_TRY(cudaGetLastError());
    DAFS_CUDA_TRY(cudaMemcpyAsync(p->h_chunk_end + c, p->d_nz_off + k0 + kn, sizeof(int64_t), cudaMemcpyDeviceToHost, p->stream));
    DAFS_CUDA_TRY(cudaEventRecord(p->ev_chunk[c], p->stream));
  }
  DAFS_CUDA_TRY(cudaEventRecord(p->ev1, p->stream));
  p->launched = true;
  p->nnz = -1;
  return DAFS_CUDA_OK;
}

}  // namespace dafs

// ---- C ABI -------------------------------------------------------------------
using dafs::fail;

extern "C" int dafs_cuda_pairhmm_tables(float* out54) {
  if (!out54) return fail(DAFS_CUDA_EINVAL, "out54 is NULL");
  const dafs::PairHmmTables& t = dafs::host_tables();
  int k = 0;
  for (int i = 0; i < 3; ++i) out54[k++] = t.init[i];
  for (int i = 0; i < 9; ++i) out54[k++] = t.trans[i];
  for (int a = 0; a < 6; ++a) out54[k++] = t.ins[a];
  for (int a = 0; a < 6; ++a)
    for (int b = 0; b < 6; ++b) out54[k++] = t.match[a * 8 + b];
  return DAFS_CUDA_OK;
}

namespace dafs {
static thread_local bool g_chunked_fetch = false;
// Copy time of a result over kernel time of the run that produced it, as the one-call entry points of this
// process last saw it (0.47 for one B200 on a PCIe 5 link; near 1 when several ranks share the host's ingest
// bandwidth and each runs an N-th of the pairs).  Decides how a pair list is cut into launches: the copy of
// launch c runs under launch c + 1, so launch c + 1 should take about ratio x as long as launch c.
static std::atomic<float> g_copy_ratio{0.4f};
}

extern "C" int dafs_cuda_pairhmm_plan_create(int nseq, const char* const* seq, const int* len,
                                             int64_t npairs, const int32_t* pi, const int32_t* pj,
                                             float th, dafs_pairhmm_plan** out) {
  using namespace dafs;
  if (!out) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  *out = nullptr;
  if (nseq < 0 || npairs < 0 || npairs > INT32_MAX / 2) return fail(DAFS_CUDA_EINVAL, "bad nseq/npairs");
  if ((nseq > 0 && (!seq || !len)) || (npairs > 0 && (!pi || !pj))) return fail(DAFS_CUDA_EINVAL, "NULL input array");
  if (std::isnan(th)) return fail(DAFS_CUDA_EINVAL, "threshold is NaN");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  }
  for (int i = 0; i < nseq; ++i)
    if (len[i] < 1 || !seq[i]) return fail(DAFS_CUDA_EINVAL, "sequence %d is empty (the reference reads seq[1] unconditionally, ProbabilisticModel.h:123)", i);
  for (int64_t k = 0; k < npairs; ++k)
    if (pi[k] < 0 || pi[k] >= nseq || pj[k] < 0 || pj[k] >= nseq) return fail(DAFS_CUDA_EINVAL, "pair %lld out of range", (long long)k);

  dafs_pairhmm_plan* p = new dafs_pairhmm_plan();
  auto bail = [&](int rc) { free_plan_buffers(p); delete p; return rc; };
#define PLAN_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return bail(fail(e_ == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM : DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_))); } while (0)
  PLAN_TRY(cudaGetDevice(&p->device));
  p->nseq = nseq; p->npairs = (int)npairs; p->th = th; p->xcut = compute_xcut(th);
  PLAN_TRY(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
  PLAN_TRY(cudaEventCreate(&p->ev0));
  if (g_chunked_fetch) { PLAN_TRY(cudaEventCreate(&p->ev_cp0)); PLAN_TRY(cudaEventCreate(&p->ev_cp1)); }
  PLAN_TRY(cudaEventCreate(&p->ev1));
  PLAN_TRY(cudaEventCreate(&p->evk0));
  PLAN_TRY(cudaEventCreate(&p->evk1));
  PLAN_TRY(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
  for (int c = 0; c < 4; ++c) {
    PLAN_TRY(cudaEventCreateWithFlags(&p->ev_chunk[c], cudaEventDisableTiming));
    PLAN_TRY(cudaEventCreate(&p->evck0[c]));
    PLAN_TRY(cudaEventCreate(&p->evck1[c]));
  }
  p->h_chunk_end = static_cast<int64_t*>(pinned_alloc(64));
  if (!p->h_chunk_end) return bail(fail(DAFS_CUDA_ENOMEM, "pinned allocation failed"));
  {
    static std::mutex mu;
    static bool tables_up[64] = {false};
    std::lock_guard<std::mutex> g(mu);
    if (!tables_up[p->device & 63]) {
      PLAN_TRY(cudaMemcpyToSymbol(c_tab, &host_tables(), sizeof(PairHmmTables)));
      tables_up[p->device & 63] = true;
    }
  }

  // encode sequences
  std::vector<int64_t> seq_off(nseq + 1, 0);
  for (int i = 0; i < nseq; ++i) seq_off[i + 1] = seq_off[i] + len[i];
  std::vector<uint8_t> enc(std::max<int64_t>(seq_off[nseq], 1));
  for (int i = 0; i < nseq; ++i)
    for (int k = 0; k < len[i]; ++k) enc[seq_off[i] + k] = symbol_class((unsigned char)seq[i][k]);

  // pair list, cost order (longest first), row offsets, geometry
  std::vector<int2> pairs(std::max<int64_t>(npairs, 1));
  std::vector<int32_t> order(std::max<int64_t>(npairs, 1));
  p->row_off.assign(npairs + 1, 0);
  int max_l1 = 1, max_l2 = 1;
  int64_t est_nnz = 0;
  for (int64_t k = 0; k < npairs; ++k) {
    pairs[k] = make_int2(pi[k], pj[k]);
    order[k] = (int32_t)k;
    const int l1 = len[pi[k]], l2 = len[pj[k]];
    p->row_off[k + 1] = p->row_off[k] + l1 + 1;
    max_l1 = std::max(max_l1, l1); max_l2 = std::max(max_l2, l2);
    est_nnz += (int64_t)l1 * std::min(l2, 10) + 64;
  }
  p->total_rows = p->row_off[npairs];
  p->C = choose_cols(max_l2);
  const int sw = 32 * p->C;
  const int64_t pw = (int64_t)((max_l2 + 1 + sw - 1) / sw) * sw;
  p->plane_stride = (int64_t)(max_l1 + 32) * pw;     // pipeline-order plane: L1 + 32 steps per stripe
  p->bnd_stride = max_l1 + 1;
  dafs::DevInfo prop;
  PLAN_TRY(dafs::dev_info(p->device, &prop));
  const int max_grid = prop.sms * kMinBlocks;
  p->grid = (int)std::max<int64_t>(1, std::min<int64_t>(max_grid, (npairs + kWarps - 1) / kWarps));
  {
    // three planes per warp slot; long sequences get fewer resident CTAs rather
    // than more than 60 % of the free HBM
    size_t free_b = 0, total_b = 0;
    PLAN_TRY(cudaMemGetInfo(&free_b, &total_b));
    const int64_t per_cta = (int64_t)sizeof(float) * 3 * p->plane_stride * kWarps;
    const int64_t fit = (int64_t)((double)free_b * 0.6) / std::max<int64_t>(per_cta, 1);
    if (fit < 1) return bail(fail(DAFS_CUDA_ENOMEM, "pair-HMM workspace for %d x %d does not fit in HBM", max_l1, max_l2));
    p->grid = (int)std::min<int64_t>(p->grid, fit);
  }
  if (max_l1 > 65534 || max_l2 > 65534) return bail(fail(DAFS_CUDA_EINVAL, "sequences longer than 65534 are not supported"));
  const int64_t slots = (int64_t)p->grid * kWarps;
  // chunks: whole rounds of the resident warps, the last (whose copy is not hidden) the smallest;
  // inside a chunk the queue hands out the most expensive pairs first
  {
    const int64_t rounds = (npairs + slots - 1) / slots;
    // Chunk boundaries cost about 4 % of device time (drain and refill of the resident
    // warps), so only the one-call entry points, whose result always goes to the host, ask
    // for them; a split-phase plan runs as one launch unless DAFS_PAIRHMM_CHUNKS says otherwise.
    const double ratio = std::min(1.0, std::max(0.25, (double)g_copy_ratio.load()));
    int want = 1;
    if (g_chunked_fetch) {
      if (rounds >= 6) want = ratio > 0.6 ? 4 : 3;
      else if (rounds >= 2 && ratio >= 0.6) want = (int)std::min<int64_t>(rounds, 4);   // the copy is as long as the kernels
    }
    if (const char* e = getenv("DAFS_PAIRHMM_CHUNKS")) want = std::max(1, std::min(4, atoi(e)));
    if (rounds < want) want = 1;
    p->nchunks = want;
    // Sizes fall geometrically (x ratio): the copy of chunk c runs under the kernel of chunk
    // c + 1, which only has to be ratio x as long (one GPU: 558 MB leave at 57 GB/s = 9.8 ms, the kernels
    // take 21 ms), and the copy of the last chunk, which nothing hides, is the smallest.
    int64_t rc[4] = {0, 0, 0, 0};
    {
      double wsum = 0.0, w = 1.0;
      for (int c = 0; c < want; ++c) { wsum += w; w *= ratio; }
      int64_t used = 0;
      for (int c = want - 1; c >= 1; --c) {                 // smallest first, the first chunk takes the rest
        double wc = 1.0;
        for (int k = 0; k < c; ++k) wc *= ratio;
        rc[c] = std::max<int64_t>(1, (int64_t)((double)rounds * wc / wsum + 0.5));
        used += rc[c];
      }
      rc[0] = rounds - used;
      if (want > 1 && rc[0] < rc[1]) {                       // few rounds: fall back to an even split
        int64_t left_rounds = rounds;
        for (int c = 0; c < want; ++c) { rc[c] = (left_rounds + (want - c) - 1) / (want - c); left_rounds -= rc[c]; }
      }
    }
    int64_t k0 = 0;
    for (int c = 0; c < want; ++c) {
      p->chunk_k0[c] = (int)k0;
      k0 = std::min<int64_t>(npairs, k0 + rc[c] * slots);
    }
    p->chunk_k0[want] = (int)npairs;
    for (int c = 0; c < want; ++c)
      std::stable_sort(order.begin() + p->chunk_k0[c], order.begin() + p->chunk_k0[c + 1], [&](int32_t a, int32_t b) {
        return (int64_t)len[pi[a]] * len[pj[a]] > (int64_t)len[pi[b]] * len[pj[b]];
      });
  }

  PLAN_TRY(dev_malloc(&p->d_seq, enc.size()));
  PLAN_TRY(dev_malloc(&p->d_seq_off, sizeof(int64_t) * (nseq + 1)));
  PLAN_TRY(dev_malloc(&p->d_seq_len, sizeof(int32_t) * std::max(nseq, 1)));
  PLAN_TRY(dev_malloc(&p->d_pairs, sizeof(int2) * pairs.size()));
  PLAN_TRY(dev_malloc(&p->d_order, sizeof(int32_t) * order.size()));
  PLAN_TRY(dev_malloc(&p->d_planes, sizeof(float) * 3 * p->plane_stride * slots));
  PLAN_TRY(dev_malloc(&p->d_bnd, sizeof(float4) * p->bnd_stride * 2 * slots));
  PLAN_TRY(dev_malloc(&p->d_queue, sizeof(int)));
  PLAN_TRY(dev_malloc(&p->d_pool_count, sizeof(unsigned long long)));
  PLAN_TRY(dev_malloc(&p->d_row_off, sizeof(int64_t) * (npairs + 1)));
  PLAN_TRY(dev_malloc(&p->d_row_ptr, sizeof(int32_t) * std::max<int64_t>(p->total_rows, 1)));
  PLAN_TRY(dev_malloc(&p->d_pair_pool_off, sizeof(int64_t) * pairs.size()));
  PLAN_TRY(dev_malloc(&p->d_pair_nnz, sizeof(int32_t) * pairs.size()));
  PLAN_TRY(dev_malloc(&p->d_pair_total, sizeof(float) * pairs.size()));
  PLAN_TRY(dev_malloc(&p->d_nz_off, sizeof(int64_t) * (npairs + 1)));
  { const int rc = alloc_pool(p, est_nnz); if (rc) return bail(rc); }
  PLAN_TRY(cudaMemcpyAsync(p->d_seq, enc.data(), enc.size(), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_seq_off, seq_off.data(), sizeof(int64_t) * (nseq + 1), cudaMemcpyHostToDevice, p->stream));
  if (nseq) PLAN_TRY(cudaMemcpyAsync(p->d_seq_len, len, sizeof(int32_t) * nseq, cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_pairs, pairs.data(), sizeof(int2) * pairs.size(), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_order, order.data(), sizeof(int32_t) * order.size(), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaMemcpyAsync(p->d_row_off, p->row_off.data(), sizeof(int64_t) * (npairs + 1), cudaMemcpyHostToDevice, p->stream));
  PLAN_TRY(cudaStreamSynchronize(p->stream));   // host vectors above go out of scope
#undef PLAN_TRY
  *out = p;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_pairhmm_plan_run(dafs_pairhmm_plan* p) {
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result arrays were handed over (dafs_cuda_pairhmm_plan_fetch_device): create a new plan");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  if (p->npairs == 0) { p->launched = true; p->nnz = 0; return DAFS_CUDA_OK; }
  return dafs::launch_all(p);
}

extern "C" int dafs_cuda_pairhmm_plan_sync(dafs_pairhmm_plan* p, float* device_ms, int64_t* nnz) {
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  if (p->npairs > 0 && p->nnz < 0) {
    for (int attempt = 0;; ++attempt) {
      DAFS_CUDA_TRY(cudaStreamSynchronize(p->stream));
      unsigned long long count = 0;
      DAFS_CUDA_TRY(cudaMemcpy(&count, p->d_pool_count, sizeof count, cudaMemcpyDeviceToHost));
      if ((int64_t)count <= p->pool_cap) { p->nnz = (int64_t)count; break; }
      // The pool was sized from a density guess; now the exact size is known.
      if (attempt >= 2) return fail(DAFS_CUDA_ECUDA, "posterior pool overflow persists");
      const int rc = dafs::alloc_pool(p, (int64_t)count + 1024);
      if (rc) return rc;
      const int rc2 = dafs::launch_all(p);
      if (rc2) return rc2;
    }
    DAFS_CUDA_TRY(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    p->last_kernel_ms = 0.0f;
    for (int c = 0; c < p->nchunks; ++c) {
      float ms = 0.0f;
      DAFS_CUDA_TRY(cudaEventElapsedTime(&ms, p->evck0[c], p->evck1[c]));
      p->last_kernel_ms += ms;
    }
  }
  if (device_ms) *device_ms = p->last_ms;
  if (nnz) *nnz = p->nnz;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_pairhmm_plan_info(dafs_pairhmm_plan* p, dafs_pairhmm_info* info) {
  if (!p || !info) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  info->cols_per_lane = p->C; info->grid = p->grid; info->warps_per_cta = dafs::kWarps;
  info->workspace_bytes = (int64_t)sizeof(float) * 3 * p->plane_stride * p->grid * dafs::kWarps;
  info->last_total_ms = p->last_ms; info->last_kernel_ms = p->last_kernel_ms;
  info->launches_per_run = 3 * p->nchunks;
  return DAFS_CUDA_OK;
}

namespace dafs {
struct HostCsr { int64_t* nz_off; int32_t* row_ptr; uint32_t* col; float* val; };

// Copies the CSR of a run that is still in flight, chunk by chunk on the copy stream, each chunk
// as soon as it is assembled (under the next chunk's kernel).  cap = entries col / val can take.
// *overflow: a chunk ended past cap (pool too small, or the caller's buffers): nothing usable was
// delivered, the caller synchronises (which regrows the pool and reruns if need be) and copies
// once with the exact size.
static cudaError_t copy_chunks(dafs_pairhmm_plan* p, const HostCsr& h, int64_t cap, bool* overflow, int64_t* copied) {
  cudaError_t e = cudaSuccess;
  int64_t beg = 0;
  *overflow = false;
  float copy_ms = 0.f;
  for (int c = 0; c < p->nchunks && e == cudaSuccess; ++c) {
    e = cudaEventSynchronize(p->ev_chunk[c]);
    if (e != cudaSuccess) break;
    if (c > 0 && p->ev_cp0) {                          // the previous chunk's copy is over or still running: either way
      float ms = 0.f;                                  // its own duration is what the events give
      if (cudaEventSynchronize(p->ev_cp1) == cudaSuccess && cudaEventElapsedTime(&ms, p->ev_cp0, p->ev_cp1) == cudaSuccess) copy_ms += ms;
    }
    if (p->ev_cp0) cudaEventRecord(p->ev_cp0, p->copy_stream);
    const int64_t end = p->h_chunk_end[c];
    if (end > cap || end > p->pool_cap) { *overflow = true; break; }
    const int64_t r0 = p->row_off[p->chunk_k0[c]], r1 = p->row_off[p->chunk_k0[c + 1]];
    if (r1 > r0) e = cudaMemcpyAsync(h.row_ptr + r0, p->d_row_ptr + r0, sizeof(int32_t) * (r1 - r0), cudaMemcpyDeviceToHost, p->copy_stream);
    if (e == cudaSuccess && end > beg) e = cudaMemcpyAsync(h.col + beg, p->d_col + beg, sizeof(uint32_t) * (end - beg), cudaMemcpyDeviceToHost, p->copy_stream);
    if (e == cudaSuccess && end > beg) e = cudaMemcpyAsync(h.val + beg, p->d_val + beg, sizeof(float) * (end - beg), cudaMemcpyDeviceToHost, p->copy_stream);
    if (p->ev_cp1) cudaEventRecord(p->ev_cp1, p->copy_stream);
    beg = end;
  }
  if (e == cudaSuccess && !*overflow) e = cudaMemcpyAsync(h.nz_off, p->d_nz_off, sizeof(int64_t) * (p->npairs + 1), cudaMemcpyDeviceToHost, p->copy_stream);
  const cudaError_t e2 = cudaStreamSynchronize(p->copy_stream);
  if (e == cudaSuccess && e2 == cudaSuccess && !*overflow && p->ev_cp0) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->ev_cp0, p->ev_cp1) == cudaSuccess) copy_ms += ms;
    p->last_copy_ms = copy_ms;
  }
  *copied = beg;
  return e == cudaSuccess ? e2 : e;
}

// One copy of a synchronised result (p->nnz known).
static cudaError_t copy_whole(dafs_pairhmm_plan* p, const HostCsr& h) {
  const int64_t nnz = p->nnz, rows = p->total_rows;
  if (p->ev_cp0) cudaEventRecord(p->ev_cp0, p->stream);
  cudaError_t e = cudaMemcpyAsync(h.nz_off, p->d_nz_off, sizeof(int64_t) * (p->npairs + 1), cudaMemcpyDeviceToHost, p->stream);
  if (e == cudaSuccess && rows) e = cudaMemcpyAsync(h.row_ptr, p->d_row_ptr, sizeof(int32_t) * rows, cudaMemcpyDeviceToHost, p->stream);
  if (e == cudaSuccess && nnz) e = cudaMemcpyAsync(h.col, p->d_col, sizeof(uint32_t) * nnz, cudaMemcpyDeviceToHost, p->stream);
  if (e == cudaSuccess && nnz) e = cudaMemcpyAsync(h.val, p->d_val, sizeof(float) * nnz, cudaMemcpyDeviceToHost, p->stream);
  if (p->ev_cp1) cudaEventRecord(p->ev_cp1, p->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
  if (e == cudaSuccess && p->ev_cp0) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->ev_cp0, p->ev_cp1) == cudaSuccess) p->last_copy_ms = ms;
  }
  return e;
}
}  // namespace dafs

extern "C" int dafs_cuda_pairhmm_plan_fetch(dafs_pairhmm_plan* p, dafs_csr_batch* out) {
  if (!p || !out) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  *out = dafs_csr_batch{};
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was handed over to a device batch");
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  const int64_t np = p->npairs, rows = p->total_rows;
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t b_row_off = sizeof(int64_t) * (np + 1), b_nz_off = sizeof(int64_t) * (np + 1);
  const size_t b_row_ptr = sizeof(int32_t) * rows;
  // One pinned block: row_off | nz_off | row_ptr | col | val.  When the run is still in flight
  // and chunked, col / val are laid out for the pool capacity (an upper bound of nnz) and every
  // chunk is copied on a second stream as soon as its CSR is assembled, under the next chunk's
  // kernel; otherwise (already synchronised, one chunk, or the pool overflowed and the run has
  // to be repeated) one copy after the exact size is known.
  for (int mode = (p->nnz < 0 && p->nchunks > 1 && np > 0) ? 0 : 1; mode < 2; ++mode) {
    int64_t cap = 0;
    if (mode == 0) cap = p->pool_cap;
    else {
      const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);
      if (rc) return rc;
      cap = p->nnz;
    }
    const size_t b_col = sizeof(uint32_t) * cap, b_val = sizeof(float) * cap;
    const size_t total = up(b_row_off) + up(b_nz_off) + up(b_row_ptr) + up(b_col) + up(b_val);
    char* base = (char*)dafs::pinned_alloc(total);
    if (!base) return fail(DAFS_CUDA_ENOMEM, "pinned allocation of %zu bytes failed", total);
    int64_t* h_row_off = (int64_t*)base;
    dafs::HostCsr h;
    h.nz_off = (int64_t*)(base + up(b_row_off));
    h.row_ptr = (int32_t*)((char*)h.nz_off + up(b_nz_off));
    h.col = (uint32_t*)((char*)h.row_ptr + up(b_row_ptr));
    h.val = (float*)((char*)h.col + up(b_col));
    std::memcpy(h_row_off, p->row_off.data(), b_row_off);
    cudaError_t e = cudaSuccess;
    if (mode == 0) {
      bool overflow = false; int64_t copied = 0;
      e = dafs::copy_chunks(p, h, cap, &overflow, &copied);
      if (e == cudaSuccess && !overflow) {
        const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);   // sets nnz and the timings; the stream is idle
        if (rc) { dafs::pinned_free(base); return rc; }
        if (p->nnz != copied) overflow = true;                               // cannot happen: nz_off ends at the pool count
      }
      if (e == cudaSuccess && overflow) { dafs::pinned_free(base); continue; }   // regrow + rerun in sync(), then one copy
    } else if (np > 0) {
      e = dafs::copy_whole(p, h);
    } else {
      h.nz_off[0] = 0;
    }
    if (e != cudaSuccess) {
      dafs::pinned_free(base);
      return fail(DAFS_CUDA_ECUDA, "result copy failed: %s", cudaGetErrorString(e));
    }
    out->n_units = np; out->nnz = p->nnz;
    out->row_off = h_row_off; out->row_ptr = h.row_ptr; out->nz_off = h.nz_off;
    out->col = h.col; out->val = h.val; out->owner_ = base;
    return DAFS_CUDA_OK;
  }
  return fail(DAFS_CUDA_ECUDA, "internal: fetch fell through");
}

// The same delivery into host arrays the CALLER owns (dafs_host_csr): what one rank of a
// multi-process run calls with its region of a host buffer that the consuming process maps
// (page-locked by the caller so that the chunk copies run under the kernels).
extern "C" int dafs_cuda_pairhmm_plan_fetch_into(dafs_pairhmm_plan* p, dafs_host_csr* dst) {
  if (!p || !dst) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was handed over to a device batch");
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  if (!dst->nz_off || !dst->row_ptr || (dst->cap_nnz > 0 && (!dst->col || !dst->val)))
    return fail(DAFS_CUDA_EINVAL, "dafs_host_csr: NULL array");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  dst->n_units = p->npairs; dst->rows = p->total_rows; dst->nnz = 0;
  if (dst->cap_units < p->npairs || dst->cap_rows < p->total_rows)
    return fail(DAFS_CUDA_ENOMEM, "dafs_host_csr holds %lld units / %lld rows, the result has %lld / %lld",
                (long long)dst->cap_units, (long long)dst->cap_rows, (long long)p->npairs, (long long)p->total_rows);
  if (p->npairs == 0) { dst->nz_off[0] = 0; return DAFS_CUDA_OK; }
  dafs::HostCsr h{dst->nz_off, dst->row_ptr, dst->col, dst->val};
  if (p->nnz < 0 && p->nchunks > 1) {
    bool overflow = false; int64_t copied = 0;
    const cudaError_t e = dafs::copy_chunks(p, h, dst->cap_nnz, &overflow, &copied);
    if (e != cudaSuccess) return fail(DAFS_CUDA_ECUDA, "result copy failed: %s", cudaGetErrorString(e));
    const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);
    if (rc) return rc;
    if (!overflow && p->nnz == copied) { dst->nnz = p->nnz; return DAFS_CUDA_OK; }
  } else {
    const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);
    if (rc) return rc;
  }
  dst->nnz = p->nnz;
  if (p->nnz > dst->cap_nnz)
    return fail(DAFS_CUDA_ENOMEM, "dafs_host_csr holds %lld entries, the result has %lld (nnz is set: regrow and call again)",
                (long long)dst->cap_nnz, (long long)p->nnz);
  const cudaError_t e = dafs::copy_whole(p, h);
  if (e != cudaSuccess) return fail(DAFS_CUDA_ECUDA, "result copy failed: %s", cudaGetErrorString(e));
  return DAFS_CUDA_OK;
}

// The result stays in HBM: row_ptr / col / val of the batch are the plan's own arrays (the plan gives
// them up and cannot be run again), row_off / nz_off are host copies.  What a device consumer
// (dafs_cuda_pct, dafs_cuda_progressive) takes instead of a host batch.
extern "C" int dafs_cuda_pairhmm_plan_fetch_device(dafs_pairhmm_plan* p, dafs_csr_batch* out) {
  if (!p || !out) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  *out = dafs_csr_batch{};
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was already handed over");
  const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);
  if (rc) return rc;
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  if (p->npairs == 0) DAFS_CUDA_TRY(cudaMemsetAsync(p->d_nz_off, 0, sizeof(int64_t), p->stream));
  const int rc2 = dafs::make_device_batch(p->npairs, p->nnz, p->row_off.data(), p->d_nz_off, p->d_row_ptr, p->d_col, p->d_val, p->stream, out);
  if (rc2) return rc2;
  p->d_row_ptr = nullptr; p->d_col = nullptr; p->d_val = nullptr;
  p->consumed = true;
  return DAFS_CUDA_OK;
}

// Device-to-device copy of the pair-ordered CSR into caller-owned device
// buffers (e.g. tensors of the process that will run an NCCL all-gather on them).
extern "C" int dafs_cuda_pairhmm_plan_export(dafs_pairhmm_plan* p, int32_t* d_row_ptr,
                                             int64_t* d_nz_off, uint32_t* d_col, float* d_val,
                                             int64_t cap_nnz) {
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  if (p->consumed) return fail(DAFS_CUDA_EINVAL, "the plan's result was handed over to a device batch");
  const int rc = dafs_cuda_pairhmm_plan_sync(p, nullptr, nullptr);
  if (rc) return rc;
  if (p->nnz > cap_nnz) return fail(DAFS_CUDA_EINVAL, "export buffers hold %lld entries, result has %lld", (long long)cap_nnz, (long long)p->nnz);
  if (p->npairs == 0) return DAFS_CUDA_OK;
  if (d_row_ptr) DAFS_CUDA_TRY(cudaMemcpyAsync(d_row_ptr, p->d_row_ptr, sizeof(int32_t) * p->total_rows, cudaMemcpyDeviceToDevice, p->stream));
  if (d_nz_off) DAFS_CUDA_TRY(cudaMemcpyAsync(d_nz_off, p->d_nz_off, sizeof(int64_t) * (p->npairs + 1), cudaMemcpyDeviceToDevice, p->stream));
  if (d_col && p->nnz) DAFS_CUDA_TRY(cudaMemcpyAsync(d_col, p->d_col, sizeof(uint32_t) * p->nnz, cudaMemcpyDeviceToDevice, p->stream));
  if (d_val && p->nnz) DAFS_CUDA_TRY(cudaMemcpyAsync(d_val, p->d_val, sizeof(float) * p->nnz, cudaMemcpyDeviceToDevice, p->stream));
  DAFS_CUDA_TRY(cudaStreamSynchronize(p->stream));
  return DAFS_CUDA_OK;
}

extern "C" void dafs_cuda_pairhmm_plan_destroy(dafs_pairhmm_plan* p) {
  if (!p) return;
  cudaSetDevice(p->device);
  if (p->stream) cudaStreamSynchronize(p->stream);
  dafs::free_plan_buffers(p);
  delete p;
}

namespace dafs {
// after a one-call run: remember how long its result took to copy relative to its kernels
static void note_copy_ratio(const dafs_pairhmm_plan* p) {
  if (p->last_copy_ms > 0.f && p->last_kernel_ms > 0.05f && p->nnz > (1 << 20)) {
    const float r = p->last_copy_ms / p->last_kernel_ms;
    g_copy_ratio.store(0.5f * g_copy_ratio.load() + 0.5f * std::min(4.0f, r));
  }
}
}  // namespace dafs

extern "C" int dafs_cuda_pairhmm_pairs(int nseq, const char* const* seq, const int* len,
                                       int64_t npairs, const int32_t* pi, const int32_t* pj,
                                       float th, dafs_csr_batch* out) {
  if (!out) return fail(DAFS_CUDA_EINVAL, "out is NULL");
  *out = dafs_csr_batch{};
  dafs_pairhmm_plan* p = nullptr;
  dafs::g_chunked_fetch = true;     // the result goes to the host: copy chunk by chunk under the kernel
  int rc = dafs_cuda_pairhmm_plan_create(nseq, seq, len, npairs, pi, pj, th, &p);
  dafs::g_chunked_fetch = false;
  if (rc) return rc;
  rc = dafs_cuda_pairhmm_plan_run(p);
  if (!rc) rc = dafs_cuda_pairhmm_plan_fetch(p, out);
  if (!rc) dafs::note_copy_ratio(p);
  dafs_cuda_pairhmm_plan_destroy(p);
  return rc;
}

extern "C" int dafs_cuda_pairhmm_pairs_into(int nseq, const char* const* seq, const int* len,
                                            int64_t npairs, const int32_t* pi, const int32_t* pj,
                                            float th, dafs_host_csr* dst) {
  if (!dst) return fail(DAFS_CUDA_EINVAL, "dst is NULL");
  dafs_pairhmm_plan* p = nullptr;
  dafs::g_chunked_fetch = true;
  int rc = dafs_cuda_pairhmm_plan_create(nseq, seq, len, npairs, pi, pj, th, &p);
  dafs::g_chunked_fetch = false;
  if (rc) return rc;
  rc = dafs_cuda_pairhmm_plan_run(p);
  if (!rc) rc = dafs_cuda_pairhmm_plan_fetch_into(p, dst);
  if (!rc) dafs::note_copy_ratio(p);
  dafs_cuda_pairhmm_plan_destroy(p);
  return rc;
}

// All i<j pairs, row-major (src/align.cpp:41-51).  With ngpu > 1 the pair list
// is dealt round-robin over the devices of this process (pairs arrive cost
// sorted inside each plan), one host thread per device, and the per-device
// batches are merged on the host in pair order.
extern "C" int dafs_cuda_pairhmm_all(int nseq, const char* const* seq, const int* len, float th,
                                     int ngpu, dafs_csr_batch* out) {
  if (!out) return fail(DAFS_CUDA_EINVAL, "out is NULL");
  *out = dafs_csr_batch{};
  if (nseq < 0) return fail(DAFS_CUDA_EINVAL, "nseq < 0");
  const int64_t npairs = (int64_t)nseq * (nseq - 1) / 2;
  std::vector<int32_t> pi(std::max<int64_t>(npairs, 1)), pj(std::max<int64_t>(npairs, 1));
  int64_t k = 0;
  for (int i = 0; i < nseq; ++i)
    for (int j = i + 1; j < nseq; ++j) { pi[k] = i; pj[k] = j; ++k; }
  const int ndev = dafs_cuda_device_count();
  if (ndev <= 0) return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  if (ngpu > ndev) return fail(DAFS_CUDA_ENODEV, "requested %d GPUs, %d visible", ngpu, ndev);
  if (ngpu <= 1 || npairs < 2 * ngpu)
    return dafs_cuda_pairhmm_pairs(nseq, seq, len, npairs, pi.data(), pj.data(), th, out);

  std::vector<std::vector<int32_t>> spi(ngpu), spj(ngpu);
  std::vector<std::vector<int64_t>> sidx(ngpu);
  for (int64_t q = 0; q < npairs; ++q) {
    const int d = (int)(q % ngpu);
    spi[d].push_back(pi[q]); spj[d].push_back(pj[q]); sidx[d].push_back(q);
  }
  std::vector<dafs_csr_batch> part(ngpu);
  std::vector<int> rc(ngpu, 0);
  std::vector<std::string> err(ngpu);
  std::vector<std::thread> workers;
  for (int d = 0; d < ngpu; ++d)
    workers.emplace_back([&, d]() {
      if (cudaSetDevice(d) != cudaSuccess) { rc[d] = DAFS_CUDA_ECUDA; err[d] = "cudaSetDevice failed"; return; }
      rc[d] = dafs_cuda_pairhmm_pairs(nseq, seq, len, (int64_t)spi[d].size(), spi[d].data(),
                                      spj[d].data(), th, &part[d]);
      if (rc[d]) err[d] = dafs::last_error();
    });
  for (auto& w : workers) w.join();
  for (int d = 0; d < ngpu; ++d)
    if (rc[d]) {
      for (auto& b : part) dafs_cuda_free_batch(&b);
      return fail(rc[d], "device %d: %s", d, err[d].c_str());
    }
  // merge
  int64_t nnz = 0, rows = 0;
  for (int64_t q = 0; q < npairs; ++q) rows += len[pi[q]] + 1;
  for (auto& b : part) nnz += b.nnz;
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t b_off = sizeof(int64_t) * (npairs + 1);
  const size_t total = 2 * up(b_off) + up(sizeof(int32_t) * rows) + up(sizeof(uint32_t) * nnz) + up(sizeof(float) * nnz);
  char* base = (char*)dafs::pinned_alloc(total);
  if (!base) { for (auto& b : part) dafs_cuda_free_batch(&b); return fail(DAFS_CUDA_ENOMEM, "pinned allocation failed"); }
  int64_t* row_off = (int64_t*)base;
  int64_t* nz_off = (int64_t*)(base + up(b_off));
  int32_t* row_ptr = (int32_t*)((char*)nz_off + up(b_off));
  uint32_t* col = (uint32_t*)((char*)row_ptr + up(sizeof(int32_t) * rows));
  float* val = (float*)((char*)col + up(sizeof(uint32_t) * nnz));
  row_off[0] = 0; nz_off[0] = 0;
  for (int64_t q = 0; q < npairs; ++q) {
    const int d = (int)(q % ngpu); const int64_t u = q / ngpu;
    const dafs_csr_batch& b = part[d];
    const int64_t nr = b.row_off[u + 1] - b.row_off[u], nz = b.nz_off[u + 1] - b.nz_off[u];
    std::memcpy(row_ptr + row_off[q], b.row_ptr + b.row_off[u], sizeof(int32_t) * nr);
    std::memcpy(col + nz_off[q], b.col + b.nz_off[u], sizeof(uint32_t) * nz);
    std::memcpy(val + nz_off[q], b.val + b.nz_off[u], sizeof(float) * nz);
    row_off[q + 1] = row_off[q] + nr; nz_off[q + 1] = nz_off[q] + nz;
  }
  for (auto& b : part) dafs_cuda_free_batch(&b);
  out->n_units = npairs; out->nnz = nnz; out->row_off = row_off; out->row_ptr = row_ptr;
  out->nz_off = nz_off; out->col = col; out->val = val; out->owner_ = base;
  return DAFS_CUDA_OK;
}
