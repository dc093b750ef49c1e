// Host side of dafs_cuda_progressive: DAFS::align(ss, aln, root) (src/dafs.cpp:1518-1537)
// with every guide-tree node whose children are finished solved in the same WAVE:
// one launch each of K11 / K12 (prog_kernel.cuh) builds p_x, p_y, p_z of all nodes of the
// wave, one batched DD plan (dd.cu) solves them, project_alignment (:766-826) merges.
// Sibling subtrees share no state (no RNG unless -r > 0), so the results are those of the
// reference's depth-first recursion.  Also dafs_cuda_guide_tree (build_tree, :447-495).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <queue>
#include <thread>
#include <vector>

#include "common.cuh"
#include "prog_kernel.cuh"

using dafs::fail;

namespace {

struct DevBufs {
  std::vector<void*> p;
  cudaStream_t stream = nullptr;        // kernels of the call run here: wait for them before the buffers go back to the cache (error returns)
  ~DevBufs() { if (stream) cudaStreamSynchronize(stream); for (void* q : p) dafs::dev_free(q); }
  template <class T> cudaError_t get(T** out, size_t count) {
    void* q = nullptr;
    const cudaError_t e = dafs::dev_malloc(&q, std::max<size_t>(count * sizeof(T), 256));
    if (e == cudaSuccess) p.push_back(q);
    *out = static_cast<T*>(q);
    return e;
  }
};

// One grow-only device block per dafs_cuda_progressive call, carved anew for every batch: the
// batch's tables and dense averages are dead once its averages are on the host, and a
// cudaMalloc in the middle of the run waits for the DD batches in flight (19 ms measured).
struct DevArena {
  char* base = nullptr;
  size_t cap = 0, used = 0;
  ~DevArena() { if (base) { cudaDeviceSynchronize(); dafs::dev_free(base); } }   // batches on other streams may still read it on an error return
  static size_t rounded(size_t bytes) { return (std::max<size_t>(bytes, 256) + 255) & ~size_t(255); }
  cudaError_t reserve(size_t bytes) {
    used = 0;
    if (bytes <= cap) return cudaSuccess;
    if (base) { dafs::dev_free(base); base = nullptr; cap = 0; }
    const size_t want = bytes + bytes / 2;                   // the first batch is usually the largest
    void* q = nullptr;
    const cudaError_t e = dafs::dev_malloc(&q, want);
    if (e == cudaSuccess) { base = static_cast<char*>(q); cap = want; }
    return e;
  }
  template <class T> cudaError_t get(T** out, size_t count) {
    const size_t b = rounded(count * sizeof(T));
    if (used + b > cap) { *out = nullptr; return cudaErrorMemoryAllocation; }
    *out = reinterpret_cast<T*>(base + used);
    used += b;
    return cudaSuccess;
  }
};

struct StreamGuard {
  cudaStream_t s = nullptr;
  cudaEvent_t a = nullptr, b = nullptr;
  ~StreamGuard() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); if (s) cudaStreamDestroy(s); }
};

struct PinGuard {
  void* p = nullptr;
  ~PinGuard() { if (p) dafs::pinned_free(p); }
};

struct PlanGuard {
  dafs_dd_plan* p = nullptr;
  ~PlanGuard() { if (p) dafs_cuda_dd_plan_destroy(p); }
};

// a sub-alignment: member sequences in the reference's row order, gap pattern per row
struct Aln {
  int L = 0;
  std::vector<int> seq;
  std::vector<uint8_t> bits;   // [rows * L], 1 = residue
  bool built = false;
};

const uint32_t kNone = 0xffffffffu;

// project_alignment, src/dafs.cpp:766-826
void merge(Aln& out, const Aln& a1, const Aln& a2, const uint32_t* z) {
  const int L1 = a1.L, L2 = a2.L;
  int matched = 0;
  for (int i = 0; i < L1; ++i) matched += z[i] != kNone;
  const int L = L1 + L2 - matched;
  // column of the merged alignment for every column of a1 / a2
  std::vector<int> c1(L1), c2(L2);
  {
    int r = 0, k = 0;
    for (int i = 0; i < L1; ++i) {
      if (z[i] != kNone) { while (k < (int)z[i]) c2[k++] = r++; c1[i] = r; c2[k++] = r++; }
      else c1[i] = r++;
    }
    while (k < L2) c2[k++] = r++;
  }
  out.L = L;
  out.seq = a1.seq;
  out.seq.insert(out.seq.end(), a2.seq.begin(), a2.seq.end());
  out.bits.assign(out.seq.size() * (size_t)L, 0);
  for (size_t q = 0; q < a1.seq.size(); ++q)
    for (int i = 0; i < L1; ++i) out.bits[q * L + c1[i]] = a1.bits[q * L1 + i];
  for (size_t q = 0; q < a2.seq.size(); ++q)
    for (int k = 0; k < L2; ++k) out.bits[(a1.seq.size() + q) * L + c2[k]] = a2.bits[q * L2 + k];
  out.built = true;
}

// project_secondary_structure (:828-876) + the consensus of align_alignments (:941-945)
void consensus(std::vector<uint32_t>& ss, int L, const uint32_t* x, int L1, const uint32_t* y, int L2, const uint32_t* z) {
  std::vector<int> c1(L1), c2(L2);
  int r = 0, k = 0;
  for (int i = 0; i < L1; ++i) {
    if (z[i] != kNone) { while (k < (int)z[i]) c2[k++] = r++; c1[i] = r; c2[k++] = r++; }
    else c1[i] = r++;
  }
  while (k < L2) c2[k++] = r++;
  std::vector<uint32_t> xx(L, kNone), yy(L, kNone);
  for (int i = 0; i < L1; ++i) if (x[i] != kNone) xx[c1[i]] = (uint32_t)c1[x[i]];
  for (int q = 0; q < L2; ++q) if (y[q] != kNone) yy[c2[q]] = (uint32_t)c2[y[q]];
  ss.assign(L, kNone);
  for (int i = 0; i < L; ++i) if (xx[i] == yy[i]) ss[i] = xx[i];
}

int grid_for(int64_t warps_wanted, int warps_per_block) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t blocks = (warps_wanted + warps_per_block - 1) / warps_per_block;
  return (int)std::max<int64_t>(1, std::min<int64_t>(blocks, (int64_t)sms * 8));
}

struct ResultOwner {
  std::vector<int32_t> seq, node_iters, node_L;
  std::vector<uint8_t> bits;
  std::vector<uint32_t> ss, node_violated;
  std::vector<float> node_score;
};

double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

// build_tree, src/dafs.cpp:447-495.  The queue holds (value, (i, j)) ordered like std::pair;
// popping the maximum whose two labels are both live merges their clusters into node n; the
// merged cluster keeps matrix slot l, its distance to a live cluster ii is
// (d[ii][l] + d[ii][r]) * value / 2 in float.
extern "C" int dafs_cuda_guide_tree(int nseq, const float* sim, int32_t* left, int32_t* right, float* score) {
  if (nseq < 1 || !sim || !left || !right) return fail(DAFS_CUDA_EINVAL, "NULL argument or nseq < 1");
  typedef std::pair<float, std::pair<uint32_t, uint32_t>> Ent;
  const uint32_t N = (uint32_t)nseq;
  uint32_t n = N;
  for (uint32_t k = 0; k < 2 * N - 1; ++k) { left[k] = right[k] = -1; if (score) score[k] = 0.0f; }
  std::vector<float> d((size_t)N * N, 0.0f);
  std::vector<uint32_t> slot(2 * N - 1, kNone);
  for (uint32_t i = 0; i < N; ++i) slot[i] = i;
  std::priority_queue<Ent> pq;
  for (uint32_t i = 0; i + 1 < N; ++i)
    for (uint32_t j = i + 1; j < N; ++j) {
      d[(size_t)i * N + j] = d[(size_t)j * N + i] = sim[(size_t)i * N + j];
      pq.push(Ent(sim[(size_t)i * N + j], std::make_pair(i, j)));
    }
  while (!pq.empty()) {
    const Ent t = pq.top();
    pq.pop();
    const uint32_t a = t.second.first, b = t.second.second;
    if (slot[a] == kNone || slot[b] == kNone) continue;
    const uint32_t l = slot[a], r = slot[b];
    slot[a] = slot[b] = kNone;
    for (uint32_t i = 0; i < n; ++i) {
      if (slot[i] == kNone) continue;
      const uint32_t ii = slot[i];
      const float v = (d[(size_t)ii * N + l] + d[(size_t)ii * N + r]) * t.first / 2;
      d[(size_t)ii * N + l] = d[(size_t)l * N + ii] = v;
      pq.push(Ent(v, std::make_pair(i, n)));
    }
    left[n] = (int32_t)a; right[n] = (int32_t)b;
    if (score) score[n] = t.first;
    slot[n++] = l;
  }
  if (n != 2 * N - 1) return fail(DAFS_CUDA_EINVAL, "similarity matrix did not yield a full tree (NaN?)");
  return DAFS_CUDA_OK;
}

extern "C" void dafs_cuda_free_prog_result(dafs_prog_result* r) {
  if (!r) return;
  delete static_cast<ResultOwner*>(r->owner_);
  *r = dafs_prog_result{};
}

extern "C" int dafs_cuda_progressive(int nseq, const int* len, const dafs_csr_batch* mp, const dafs_csr_batch* bp,
                                     const int32_t* left, const int32_t* right, const dafs_prog_params* prm,
                                     dafs_prog_result* out) {
  using namespace dafs;
  if (out) *out = dafs_prog_result{};
  if (nseq < 1 || !len || !mp || !bp || !left || !right || !prm || !out) return fail(DAFS_CUDA_EINVAL, "NULL argument or nseq < 1");
  const int n = nseq, nn = 2 * n - 1;
  const int64_t npairs = (int64_t)n * (n - 1) / 2;
  if (mp->n_units != npairs) return fail(DAFS_CUDA_EINVAL, "mp holds %lld units, %lld pairs expected", (long long)mp->n_units, (long long)npairs);
  if (bp->n_units != n) return fail(DAFS_CUDA_EINVAL, "bp holds %lld units, %d sequences expected", (long long)bp->n_units, n);
  for (int i = 0; i < n; ++i)
    if (len[i] < 1 || len[i] > 16384) return fail(DAFS_CUDA_EINVAL, "len[%d] = %d (1 .. 16384)", i, len[i]);
  // the tree: leaves 0..n-1, internal nodes n..2n-2 whose children come earlier, root last
  std::vector<int> level(nn, 0);
  {
    std::vector<char> used(nn, 0);
    for (int k = 0; k < nn; ++k) {
      if (k < n) { if (left[k] != -1 || right[k] != -1) return fail(DAFS_CUDA_EINVAL, "tree node %d is a leaf but has children", k); continue; }
      const int a = left[k], b = right[k];
      if (a < 0 || b < 0 || a >= k || b >= k || a == b || used[a] || used[b])
        return fail(DAFS_CUDA_EINVAL, "tree node %d: children (%d, %d) are not two unused earlier nodes", k, a, b);
      used[a] = used[b] = 1;
      level[k] = 1 + std::max(level[a], level[b]);
    }
  }
  // geometry of the CSR batches
  std::vector<int2> pairs((size_t)npairs);
  std::vector<int> ncols_mp((size_t)npairs);
  std::vector<int64_t> u_row_off(npairs + 1), t_row_off(npairs + 1), b_row_off(n + 1);
  {
    int64_t u = 0, ru = 0, rt = 0;
    for (int a = 0; a < n; ++a)
      for (int b = a + 1; b < n; ++b, ++u) {
        pairs[u] = make_int2(a, b); ncols_mp[u] = len[b];
        u_row_off[u] = ru; t_row_off[u] = rt;
        ru += len[a] + 1; rt += len[b] + 1;
      }
    u_row_off[npairs] = ru; t_row_off[npairs] = rt;
    int64_t rb = 0;
    for (int a = 0; a < n; ++a) { b_row_off[a] = rb; rb += len[a] + 1; }
    b_row_off[n] = rb;
  }
  for (int64_t u = 0; u <= npairs; ++u)
    if (mp->row_off[u] != u_row_off[u]) return fail(DAFS_CUDA_EINVAL, "mp->row_off[%lld] does not match the sequence lengths", (long long)u);
  for (int a = 0; a <= n; ++a)
    if (bp->row_off[a] != b_row_off[a]) return fail(DAFS_CUDA_EINVAL, "bp->row_off[%d] does not match the sequence lengths", a);
  const int64_t mp_nnz = npairs ? mp->nz_off[npairs] : 0, bp_nnz = bp->nz_off[n];
  if (mp_nnz < 0 || bp_nnz < 0 || mp_nnz != mp->nnz || bp_nnz != bp->nnz) return fail(DAFS_CUDA_EINVAL, "nnz does not match nz_off");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  }

  const double t_begin = now_ms();
  StreamGuard G;
  DAFS_CUDA_TRY(cudaStreamCreateWithFlags(&G.s, cudaStreamNonBlocking));
  DAFS_CUDA_TRY(cudaEventCreate(&G.a));
  DAFS_CUDA_TRY(cudaEventCreate(&G.b));
  cudaStream_t st = G.s;
  DevBufs B;
  B.stream = st;
  ProgArgs A{};
  A.n = n;
  int* d_len = nullptr; int2* d_pairs = nullptr; int* d_ncols = nullptr; int* d_bad = nullptr; int* d_fill = nullptr;
  int64_t *d_u_row_off = nullptr, *d_t_row_off = nullptr, *d_nz_off = nullptr, *d_b_row_off = nullptr, *d_b_nz_off = nullptr;
  int* d_rp = nullptr; uint32_t* d_col = nullptr; float* d_val = nullptr;
  int* d_b_rp = nullptr; uint32_t* d_b_col = nullptr; float* d_b_val = nullptr;
  ProgDesc* d_desc = nullptr;
  const int64_t RU = u_row_off[npairs], RT = t_row_off[npairs];
  DAFS_CUDA_TRY(B.get(&d_len, n));
  DAFS_CUDA_TRY(B.get(&d_pairs, npairs));
  DAFS_CUDA_TRY(B.get(&d_ncols, npairs));
  DAFS_CUDA_TRY(B.get(&d_bad, 1));
  DAFS_CUDA_TRY(B.get(&d_fill, RT));
  DAFS_CUDA_TRY(B.get(&d_u_row_off, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_t_row_off, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_nz_off, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_b_row_off, n + 1));
  DAFS_CUDA_TRY(B.get(&d_b_nz_off, n + 1));
  DAFS_CUDA_TRY(B.get(&d_rp, RU + RT));
  DAFS_CUDA_TRY(B.get(&d_col, 2 * mp_nnz));
  DAFS_CUDA_TRY(B.get(&d_val, 2 * mp_nnz));
  DAFS_CUDA_TRY(B.get(&d_b_rp, b_row_off[n]));
  DAFS_CUDA_TRY(B.get(&d_b_col, bp_nnz));
  DAFS_CUDA_TRY(B.get(&d_b_val, bp_nnz));
  DAFS_CUDA_TRY(B.get(&d_desc, (size_t)n * n));
  auto h2d = [&](void* d, const void* h, size_t bytes) {
    return bytes ? cudaMemcpyAsync(d, h, bytes, cudaMemcpyDefault, st) : cudaSuccess;   // mp / bp may be device-resident batches
  };
  DAFS_CUDA_TRY(h2d(d_len, len, sizeof(int) * n));
  DAFS_CUDA_TRY(h2d(d_pairs, pairs.data(), sizeof(int2) * npairs));
  DAFS_CUDA_TRY(h2d(d_ncols, ncols_mp.data(), sizeof(int) * npairs));
  DAFS_CUDA_TRY(h2d(d_u_row_off, u_row_off.data(), sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_t_row_off, t_row_off.data(), sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_nz_off, mp->nz_off, sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_b_row_off, b_row_off.data(), sizeof(int64_t) * (n + 1)));
  DAFS_CUDA_TRY(h2d(d_b_nz_off, bp->nz_off, sizeof(int64_t) * (n + 1)));
  DAFS_CUDA_TRY(h2d(d_rp, mp->row_ptr, sizeof(int) * RU));
  DAFS_CUDA_TRY(h2d(d_col, mp->col, sizeof(uint32_t) * mp_nnz));
  DAFS_CUDA_TRY(h2d(d_val, mp->val, sizeof(float) * mp_nnz));
  DAFS_CUDA_TRY(h2d(d_b_rp, bp->row_ptr, sizeof(int) * b_row_off[n]));
  DAFS_CUDA_TRY(h2d(d_b_col, bp->col, sizeof(uint32_t) * bp_nnz));
  DAFS_CUDA_TRY(h2d(d_b_val, bp->val, sizeof(float) * bp_nnz));
  DAFS_CUDA_TRY(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
  if (npairs) prog_validate_kernel<<<grid_for(npairs, 4), 128, 0, st>>>(npairs, d_u_row_off, d_nz_off, d_rp, d_col, d_ncols, d_bad);
  prog_validate_kernel<<<grid_for(n, 4), 128, 0, st>>>(n, d_b_row_off, d_b_nz_off, d_b_rp, d_b_col, d_len, d_bad);
  int bad = 0;
  DAFS_CUDA_TRY(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
  DAFS_CUDA_TRY(cudaStreamSynchronize(st));
  if (bad) return fail(DAFS_CUDA_EINVAL, "mp / bp is not a well-formed row-sparse batch (row_ptr, ascending columns, column range)");
  int launches = 2;
  if (npairs) {
    prog_transpose_kernel<<<grid_for(npairs, 4), 128, 0, st>>>(n, d_len, npairs, d_pairs, d_u_row_off, d_t_row_off, d_nz_off, d_rp,
                                                               d_col, d_val, d_rp + RU, d_col + mp_nnz, d_val + mp_nnz, d_fill);
    ++launches;
  }
  prog_desc_kernel<<<(n * n + 255) / 256, 256, 0, st>>>(n, d_u_row_off, d_t_row_off, d_nz_off, RU, mp_nnz, d_desc);
  ++launches;
  DAFS_CUDA_TRY(cudaGetLastError());
  A.len = d_len; A.rp_all = d_rp; A.col_all = d_col; A.val_all = d_val; A.desc = d_desc;
  A.b_row_off = d_b_row_off; A.b_nz_off = d_b_nz_off; A.b_row_ptr = d_b_rp; A.b_col = d_b_col; A.b_val = d_b_val;

  // ---- batches -----------------------------------------------------------------------
  // A node is ready when both children are built.  All nodes that are ready at the same time
  // go out as one batch (two averaging launches + one DD launch on the batch's own stream);
  // finished problems are harvested one by one with dafs_cuda_dd_plan_poll while their batch
  // is still running, so a node that needs 600 iterations holds up its own ancestors only.
  std::vector<Aln> aln(nn);
  for (int s = 0; s < n; ++s) {             // leaves, :1520-1526
    aln[s].L = len[s]; aln[s].seq.assign(1, s); aln[s].bits.assign(len[s], 1); aln[s].built = true;
  }
  ResultOwner* R = new ResultOwner();
  struct OwnerGuard { ResultOwner*& r; ~OwnerGuard() { delete r; } } og{R};
  R->node_iters.assign(nn, 0); R->node_L.assign(nn, 0); R->node_violated.assign(nn, 0); R->node_score.assign(nn, 0.0f);
  for (int s = 0; s < n; ++s) R->node_L[s] = len[s];
  const int max_level = level[nn - 1];
  std::vector<int> parent(nn, -1);
  for (int k = n; k < nn; ++k) { parent[left[k]] = k; parent[right[k]] = k; }
  struct Batch {
    std::vector<int> node;
    std::vector<ProgSide> sides;
    std::vector<dafs_dd_problem> prob;
    std::vector<dafs_dd_result> res;
    std::vector<std::vector<uint32_t>> xs, ys, zs;
    std::vector<int32_t> done;
    PlanGuard plan;
    cudaStream_t stream = nullptr;
    int remaining = 0;
    double t0 = 0;
  };
  // a small pool of streams for the batches: an idle one is reused, so the number of streams
  // (and of hardware queues they map to) stays at the number of batches in flight
  struct StreamPool {
    std::vector<cudaStream_t> all, idle;
    cudaStream_t poll = nullptr;
    ~StreamPool() { for (cudaStream_t s : all) cudaStreamDestroy(s); if (poll) cudaStreamDestroy(poll); g_dd_stream = nullptr; g_dd_poll_stream = nullptr; }
  } pool;
  DAFS_CUDA_TRY(cudaStreamCreateWithFlags(&pool.poll, cudaStreamNonBlocking));
  std::vector<std::unique_ptr<Batch>> inflight;
  std::vector<int> ready;
  for (int k = n; k < nn; ++k) if (left[k] < n && right[k] < n) ready.push_back(k);
  const bool serial = getenv("DAFS_PROG_WAVES") != nullptr;   // diagnostics: wait for a whole batch, like the first version
  float ms_avg = 0.0f, ms_dd = 0.0f;
  double ms_dd_host = 0.0;
  int64_t total_iters = 0;
  float root_score = 0.0f;
  int batches = 0, built = 0;
  const bool verbose = getenv("DAFS_PROG_VERBOSE") != nullptr;
  const bool host_pack = [] { const char* e = getenv("DAFS_PROG_HOST_PACK"); return e && atoi(e) > 0; }();
  double tm_sum[12] = {0}, tm_max[12] = {0};   // host phases: prep, averaging, plan create, plan run, poll, finish
  auto lap = [&](int k, double t_from) { const double d = now_ms() - t_from; tm_sum[k] += d; tm_max[k] = std::max(tm_max[k], d); return now_ms(); };
  // a solved node: merge the children (project_alignment), record, maybe release the parent
  auto finish = [&](Batch& Bt, int q) -> int {
    const int k = Bt.node[q];
    const dafs_dd_result& r = Bt.res[q];
    if (r.status != 0) return fail(DAFS_CUDA_ECUDA, "node %d: the alignment traceback left the envelope (the reference would not terminate)", k);
    const Aln& a1 = aln[left[k]];
    const Aln& a2 = aln[right[k]];
    merge(aln[k], a1, a2, Bt.zs[q].data());
    R->node_iters[k] = r.iters; R->node_violated[k] = r.violated; R->node_score[k] = r.score;
    R->node_L[k] = aln[k].L;
    total_iters += r.iters;
    if (k == nn - 1) {
      consensus(R->ss, aln[k].L, Bt.xs[q].data(), a1.L, Bt.ys[q].data(), a2.L, Bt.zs[q].data());
      root_score = r.score;
    }
    std::vector<uint8_t>().swap(aln[left[k]].bits);       // children are no longer needed
    std::vector<uint8_t>().swap(aln[right[k]].bits);
    ++built;
    const int pk = parent[k];
    if (pk >= 0) {
      const int sib = left[pk] == k ? right[pk] : left[pk];
      if (aln[sib].built) ready.push_back(pk);
    }
    return DAFS_CUDA_OK;
  };
  PinGuard dense_pin;
  size_t dense_cap = 0;
  DevArena arena;
  struct DevOne { void* p = nullptr; ~DevOne() { if (p) { cudaDeviceSynchronize(); dafs::dev_free(p); } } } pack_buf;   // scratch of the pack kernels, lent to every batch's plan
  size_t pack_cap = 0;
  while (built < n - 1) {
    if (!ready.empty()) {
      double tl = now_ms();
      std::unique_ptr<Batch> Bp(new Batch());
      Batch& Bt = *Bp;
      Bt.node.swap(ready);
      std::sort(Bt.node.begin(), Bt.node.end());
      const int W = (int)Bt.node.size();
      // sides, members and maps of the batch
      Bt.sides.resize((size_t)2 * W);
      std::vector<ProgSide>& sides = Bt.sides;
      std::vector<ProgNode> nodes(W);
      std::vector<int> mem_seq;
      std::vector<int64_t> mem_col, mem_inv;
      std::vector<int> maps;
      std::vector<int2> bp_tasks, mp_tasks;
      int64_t dense = 0;
      int accW = 32;
      for (int q = 0; q < W; ++q) {
        const int k = Bt.node[q];
        for (int sd = 0; sd < 2; ++sd) {
          const Aln& a = aln[sd == 0 ? left[k] : right[k]];
          ProgSide& S = sides[2 * q + sd];
          S.L = a.L; S.rows = (int)a.seq.size(); S.mem0 = (int)mem_seq.size(); S.pad = 0; S.out = dense;
          dense += (int64_t)a.L * a.L;
          accW = std::max(accW, (a.L + 31) & ~31);
          for (int r = 0; r < S.rows; ++r) {
            const int s = a.seq[r];
            mem_seq.push_back(s);
            mem_col.push_back((int64_t)maps.size());
            const uint8_t* g = a.bits.data() + (size_t)r * a.L;
            for (int c = 0; c < a.L; ++c) if (g[c]) maps.push_back(c);
            mem_inv.push_back((int64_t)maps.size());
            int ii = 0;
            for (int c = 0; c < a.L; ++c) maps.push_back(g[c] ? ii++ : -1);
            if (ii != len[s]) return fail(DAFS_CUDA_ECUDA, "internal: alignment row of sequence %d holds %d residues", s, ii);
          }
          for (int c = 0; c < a.L; ++c) bp_tasks.push_back(make_int2(2 * q + sd, c));
        }
        nodes[q].side1 = 2 * q; nodes[q].side2 = 2 * q + 1; nodes[q].pad0 = nodes[q].pad1 = 0; nodes[q].out = dense;
        dense += (int64_t)sides[2 * q].L * sides[2 * q + 1].L;
        for (int c = 0; c < sides[2 * q].L; ++c) mp_tasks.push_back(make_int2(q, c));
      }
      int wpb = 8;
      while (wpb > 1 && (size_t)wpb * accW * sizeof(float) > (size_t(96) << 10)) wpb /= 2;
      const size_t smem = (size_t)wpb * accW * sizeof(float);
      if (smem > (size_t(227) << 10)) return fail(DAFS_CUDA_EINVAL, "alignment of %d columns is too long for the averaging kernels", accW);
      if (smem > (48 << 10)) {
        DAFS_CUDA_TRY(cudaFuncSetAttribute(prog_avg_bp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DAFS_CUDA_TRY(cudaFuncSetAttribute(prog_avg_mp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      }
      tl = lap(0, tl);
      DevArena& WB = arena;                         // the batch's device buffers (dead after the copy below)
      DAFS_CUDA_TRY(WB.reserve(DevArena::rounded(sizeof(ProgSide) * sides.size()) + DevArena::rounded(sizeof(ProgNode) * nodes.size()) +
                               DevArena::rounded(sizeof(int) * mem_seq.size()) + DevArena::rounded(sizeof(int64_t) * mem_col.size()) +
                               DevArena::rounded(sizeof(int64_t) * mem_inv.size()) + DevArena::rounded(sizeof(int) * maps.size()) +
                               DevArena::rounded(sizeof(int2) * bp_tasks.size()) + DevArena::rounded(sizeof(int2) * mp_tasks.size()) +
                               DevArena::rounded(sizeof(float) * (size_t)dense)));
      ProgSide* d_sides = nullptr; ProgNode* d_nodes = nullptr; int* d_mem_seq = nullptr; int64_t* d_mem_col = nullptr;
      int64_t* d_mem_inv = nullptr; int* d_maps = nullptr; int2* d_bp_tasks = nullptr; int2* d_mp_tasks = nullptr; float* d_dense = nullptr;
      DAFS_CUDA_TRY(WB.get(&d_sides, sides.size()));
      DAFS_CUDA_TRY(WB.get(&d_nodes, nodes.size()));
      DAFS_CUDA_TRY(WB.get(&d_mem_seq, mem_seq.size()));
      DAFS_CUDA_TRY(WB.get(&d_mem_col, mem_col.size()));
      DAFS_CUDA_TRY(WB.get(&d_mem_inv, mem_inv.size()));
      DAFS_CUDA_TRY(WB.get(&d_maps, maps.size()));
      DAFS_CUDA_TRY(WB.get(&d_bp_tasks, bp_tasks.size()));
      DAFS_CUDA_TRY(WB.get(&d_mp_tasks, mp_tasks.size()));
      DAFS_CUDA_TRY(WB.get(&d_dense, (size_t)dense));
      tl = lap(6, tl);
      DAFS_CUDA_TRY(h2d(d_sides, sides.data(), sizeof(ProgSide) * sides.size()));
      DAFS_CUDA_TRY(h2d(d_nodes, nodes.data(), sizeof(ProgNode) * nodes.size()));
      DAFS_CUDA_TRY(h2d(d_mem_seq, mem_seq.data(), sizeof(int) * mem_seq.size()));
      DAFS_CUDA_TRY(h2d(d_mem_col, mem_col.data(), sizeof(int64_t) * mem_col.size()));
      DAFS_CUDA_TRY(h2d(d_mem_inv, mem_inv.data(), sizeof(int64_t) * mem_inv.size()));
      DAFS_CUDA_TRY(h2d(d_maps, maps.data(), sizeof(int) * maps.size()));
      DAFS_CUDA_TRY(h2d(d_bp_tasks, bp_tasks.data(), sizeof(int2) * bp_tasks.size()));
      DAFS_CUDA_TRY(h2d(d_mp_tasks, mp_tasks.data(), sizeof(int2) * mp_tasks.size()));
      A.sides = d_sides; A.nodes = d_nodes; A.mem_seq = d_mem_seq; A.mem_col = d_mem_col; A.mem_inv = d_mem_inv;
      A.maps = d_maps; A.dense = d_dense;
      DAFS_CUDA_TRY(cudaEventRecord(G.a, st));
      prog_avg_bp_kernel<<<grid_for((int64_t)bp_tasks.size(), wpb), 32 * wpb, smem, st>>>(A, (int64_t)bp_tasks.size(), d_bp_tasks, accW);
      prog_avg_mp_kernel<<<grid_for((int64_t)mp_tasks.size(), wpb), 32 * wpb, smem, st>>>(A, (int64_t)mp_tasks.size(), d_mp_tasks, accW);
      launches += 2;
      DAFS_CUDA_TRY(cudaGetLastError());
      DAFS_CUDA_TRY(cudaEventRecord(G.b, st));
      tl = lap(7, tl);
      // The dense averages stay in HBM: the batch's DD plan is built from them by the pack kernels
      // (dafs_cuda_dd_plan_create_device: CBP enumeration, envelope, packing).  DAFS_PROG_HOST_PACK=1 keeps
      // the earlier route for A/B runs: one D2H of the batch's matrices into a page-locked buffer that every
      // batch shares (cudaHostAlloc waits for running kernels, so it is sized generously by the first batch,
      // which runs with nothing in flight), set-up on the host cores.
      float* h_dense = nullptr;
      if (host_pack) {
        if (dense_cap < (size_t)dense) {
          if (dense_pin.p) { dafs::pinned_free(dense_pin.p); dense_pin.p = nullptr; }
          dense_cap = (size_t)dense + (size_t)dense / 2;
          dense_pin.p = pinned_alloc(sizeof(float) * dense_cap);
          if (!dense_pin.p) { dense_cap = 0; return fail(DAFS_CUDA_ENOMEM, "pinned allocation of %lld floats failed", (long long)dense); }
        }
        tl = lap(8, tl);
        h_dense = static_cast<float*>(dense_pin.p);
        DAFS_CUDA_TRY(cudaMemcpyAsync(h_dense, d_dense, sizeof(float) * (size_t)dense, cudaMemcpyDeviceToHost, st));
        DAFS_CUDA_TRY(cudaStreamSynchronize(st));
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, G.a, G.b);
        ms_avg += ms;
      }
      tl = lap(1, tl);
      // the batch's DD problems, one asynchronous launch (solve_by_dd, :1006-1295)
      Bt.t0 = now_ms();
      Bt.prob.resize(W); Bt.res.resize(W); Bt.xs.resize(W); Bt.ys.resize(W); Bt.zs.resize(W);
      Bt.done.assign(W, 0);
      Bt.remaining = W;
      for (int q = 0; q < W; ++q) {
        const ProgSide& S1 = sides[2 * q];
        const ProgSide& S2 = sides[2 * q + 1];
        dafs_dd_problem& P = Bt.prob[q];
        P.L1 = S1.L; P.L2 = S2.L; P.N1 = S1.rows; P.N2 = S2.rows;
        const float* base = host_pack ? h_dense : d_dense;
        P.p_x = base + S1.out; P.p_y = base + S2.out; P.p_z = base + nodes[q].out;
        P.w = prm->w; P.th_s = prm->th_s; P.min_th_s = prm->min_th_s; P.th_a = prm->th_a; P.eta0 = prm->eta0; P.t_max = prm->t_max;
        Bt.xs[q].assign(S1.L, kNone); Bt.ys[q].assign(S2.L, kNone); Bt.zs[q].assign(S1.L, kNone);
        Bt.res[q] = dafs_dd_result{};
        Bt.res[q].x = Bt.xs[q].data(); Bt.res[q].y = Bt.ys[q].data(); Bt.res[q].z = Bt.zs[q].data();
      }
      {
        if (pool.idle.empty()) {
          cudaStream_t s = nullptr;
          DAFS_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
          pool.all.push_back(s); pool.idle.push_back(s);
        }
        Bt.stream = pool.idle.back(); pool.idle.pop_back();
        g_dd_stream = Bt.stream; g_dd_poll_stream = pool.poll;
        int rc;
        if (host_pack) rc = dafs_cuda_dd_plan_create(W, Bt.prob.data(), 0, &Bt.plan.p);
        else {
          // the pack kernels run on the batch's stream, behind the averaging kernels; the call returns when
          // the blocks are built, so the arena that holds the matrices is free for the next batch
          DAFS_CUDA_TRY(cudaStreamWaitEvent(Bt.stream, G.b, 0));
          const size_t pack_need = dd_pack_bytes(W, Bt.prob.data());
          if (pack_need > pack_cap) {               // grow-only; the first batch is usually the largest and runs with nothing in flight
            dafs::dev_free(pack_buf.p); pack_buf.p = nullptr; pack_cap = 0;
            void* q = nullptr;
            DAFS_CUDA_TRY(dafs::dev_malloc(&q, pack_need + pack_need / 4));
            pack_buf.p = q; pack_cap = pack_need + pack_need / 4;
          }
          g_dd_pack_lend.base = static_cast<char*>(pack_buf.p); g_dd_pack_lend.cap = pack_cap;
          rc = dafs_cuda_dd_plan_create_device(W, Bt.prob.data(), 0, &Bt.plan.p);
          g_dd_pack_lend.base = nullptr; g_dd_pack_lend.cap = 0;
          if (!rc) { float ms = 0.0f; cudaEventElapsedTime(&ms, G.a, G.b); ms_avg += ms; }
        }
        g_dd_stream = nullptr; g_dd_poll_stream = nullptr;
        tl = lap(2, tl);
        if (!rc) rc = dafs_cuda_dd_plan_run(Bt.plan.p);
        tl = lap(3, tl);
        if (rc) return rc;
      }
      launches += 1;
      ++batches;
      inflight.push_back(std::move(Bp));
      if (!serial) continue;
    }
    if (inflight.empty()) return fail(DAFS_CUDA_ECUDA, "internal: no node is ready and none is in flight");
    // harvest
    bool progress = false;
    for (size_t bi = 0; bi < inflight.size();) {
      Batch& Bt = *inflight[bi];
      const int W = (int)Bt.node.size();
      int32_t fresh = 0;
      std::vector<int32_t> before;
      if (serial) {
        int rc = dafs_cuda_dd_plan_fetch(Bt.plan.p, Bt.res.data());
        if (rc) return rc;
        before.assign(W, 0);
        Bt.done.assign(W, 1);
        fresh = W;
      } else {
        before = Bt.done;
        const double tp = now_ms();
        const int rc = dafs_cuda_dd_plan_poll(Bt.plan.p, Bt.res.data(), Bt.done.data(), &fresh);
        lap(4, tp);
        if (rc) return rc;
      }
      if (fresh > 0) {
        progress = true;
        const double tf = now_ms();
        for (int q = 0; q < W; ++q)
          if (Bt.done[q] && !before[q]) { const int rc = finish(Bt, q); if (rc) return rc; }
        lap(5, tf);
        Bt.remaining -= fresh;
      }
      if (Bt.remaining == 0) {
        float dms = 0.0f;
        const double ts = now_ms();
        const int rc = dafs_cuda_dd_plan_sync(Bt.plan.p, &dms);
        lap(9, ts);
        if (rc) return rc;
        ms_dd += dms;
        ms_dd_host += now_ms() - Bt.t0;
        if (verbose) {                               // per-batch diagnostics for tools/bench_prog.py
          dafs_dd_info inf{};
          dafs_cuda_dd_plan_info(Bt.plan.p, &inf);
          int slow = 0;
          for (int q = 1; q < W; ++q) if (Bt.res[q].iters > Bt.res[slow].iters) slow = q;
          const float* cy = Bt.res[slow].cycles;
          const float it1 = (float)std::max(1, Bt.res[slow].iters + 1);
          fprintf(stderr, "[prog] t = %.1f .. %.1f ms: batch of %d problems (first node %d, level %d): DD %.2f ms (wall %.2f ms), grid %d x %d threads, smem %d B, in_smem %#x, band words %d, cbp %lld; "
                          "slowest: L1 %d L2 %d N %d+%d iters %d cbp %lld, cycles/iter decode %.0f update %.0f, x %.0f+%.0f y %.0f+%.0f z %.0f+%.0f\n",
                  Bt.t0 - t_begin, now_ms() - t_begin, W, Bt.node[0], level[Bt.node[0]], dms, now_ms() - Bt.t0, inf.grid, inf.block, inf.smem_bytes, inf.items_in_smem, inf.cols_per_lane, (long long)inf.total_cbp,
                  Bt.prob[slow].L1, Bt.prob[slow].L2, Bt.prob[slow].N1, Bt.prob[slow].N2, Bt.res[slow].iters, (long long)Bt.res[slow].ncbp, cy[6] / it1, cy[7] / it1,
                  cy[0] / it1, cy[1] / it1, cy[2] / it1, cy[3] / it1, cy[4] / it1, cy[5] / it1);
        }
        pool.idle.push_back(Bt.stream);
        const double te = now_ms();
        inflight.erase(inflight.begin() + bi);
        lap(10, te);
      } else {
        ++bi;
      }
    }
    if (!progress && ready.empty()) std::this_thread::sleep_for(std::chrono::microseconds(50));
  }
  if (verbose)
    fprintf(stderr, "[prog] host phases, sum / longest single (ms): prep %.1f / %.1f, averaging %.1f / %.1f, plan create %.1f / %.1f, plan run %.1f / %.1f, poll %.1f / %.1f, merge %.1f / %.1f; device alloc %.1f / %.1f, upload + launch %.1f / %.1f, pinned alloc %.1f / %.1f, plan sync %.1f / %.1f, plan destroy %.1f / %.1f\n",
            tm_sum[0], tm_max[0], tm_sum[1], tm_max[1], tm_sum[2], tm_max[2], tm_sum[3], tm_max[3], tm_sum[4], tm_max[4], tm_sum[5], tm_max[5], tm_sum[6], tm_max[6], tm_sum[7], tm_max[7], tm_sum[8], tm_max[8], tm_sum[9], tm_max[9], tm_sum[10], tm_max[10]);
  const Aln& root = aln[nn - 1];
  R->seq.assign(root.seq.begin(), root.seq.end());
  R->bits = root.bits;
  out->L = root.L; out->n_rows = n;
  out->seq = R->seq.data(); out->bits = R->bits.data();
  out->ss = R->ss.empty() ? nullptr : R->ss.data();
  out->score = root_score;
  out->n_nodes = nn;
  out->node_iters = R->node_iters.data(); out->node_violated = R->node_violated.data();
  out->node_score = R->node_score.data(); out->node_L = R->node_L.data();
  out->waves = batches; out->total_iters = total_iters; out->gpu_launches = launches;
  out->ms_average = ms_avg; out->ms_dd = ms_dd; out->ms_dd_host = (float)ms_dd_host;
  out->ms_total = (float)(now_ms() - t_begin);
  out->owner_ = R;
  R = nullptr;
  return DAFS_CUDA_OK;
}
