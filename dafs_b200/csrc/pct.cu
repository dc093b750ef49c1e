// Host side of dafs_cuda_pct: the consistency transformation of DAFS::run()
// (src/dafs.cpp:1796-1827) on the device.  Kernels: pct_kernel.cuh.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "pct_kernel.cuh"

using dafs::fail;

namespace {

struct DevBufs {                       // everything dev_malloc'ed by one call
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
  void release(void* q) {              // the caller takes the buffer over
    for (size_t k = 0; k < p.size(); ++k) if (p[k] == q) { p.erase(p.begin() + k); break; }
  }
  void drop(void* q) {
    for (size_t k = 0; k < p.size(); ++k) if (p[k] == q) { p.erase(p.begin() + k); break; }
    dafs::dev_free(q);
  }
};

thread_local float g_ms[5] = {0, 0, 0, 0, 0};
thread_local bool g_device_out = false;   // dafs_cuda_pct_device: the outputs stay in HBM

size_t up256(size_t x) { return (x + 255) & ~size_t(255); }

// pinned block for one dafs_csr_batch
struct HostBatch {
  char* base = nullptr;
  int64_t* row_off; int64_t* nz_off; int32_t* row_ptr; uint32_t* col; float* val;
  size_t b_off, b_row_ptr, b_col, b_val;
  bool alloc(int64_t units, int64_t row_entries, int64_t nnz) {
    b_off = sizeof(int64_t) * (units + 1); b_row_ptr = sizeof(int32_t) * row_entries;
    b_col = sizeof(uint32_t) * nnz; b_val = sizeof(float) * nnz;
    base = (char*)dafs::pinned_alloc(2 * up256(b_off) + up256(b_row_ptr) + up256(b_col) + up256(b_val) + 256);
    if (!base) return false;
    row_off = (int64_t*)base;
    nz_off = (int64_t*)(base + up256(b_off));
    row_ptr = (int32_t*)((char*)nz_off + up256(b_off));
    col = (uint32_t*)((char*)row_ptr + up256(b_row_ptr));
    val = (float*)((char*)col + up256(b_col));
    return true;
  }
  void give(dafs_csr_batch* out, int64_t units, int64_t nnz) {
    out->n_units = units; out->nnz = nnz; out->row_off = row_off; out->row_ptr = row_ptr; out->nz_off = nz_off;
    out->col = col; out->val = val; out->owner_ = base;
    base = nullptr;
  }
  ~HostBatch() { if (base) dafs::pinned_free(base); }
};

int copy_batch(const dafs_csr_batch* in, dafs_csr_batch* out) {
  const int64_t units = in->n_units, re = in->row_off[units], nnz = in->nz_off[units];
  if (g_device_out) {
    int32_t* d_rp = nullptr; uint32_t* d_col = nullptr; float* d_val = nullptr; int64_t* d_nz = nullptr;
    auto bail = [&](int rc) { dafs::dev_free(d_rp); dafs::dev_free(d_col); dafs::dev_free(d_val); dafs::dev_free(d_nz); return rc; };
    cudaError_t e = dafs::dev_malloc(&d_rp, std::max<size_t>(sizeof(int32_t) * re, 256));
    if (e == cudaSuccess) e = dafs::dev_malloc(&d_col, std::max<size_t>(sizeof(uint32_t) * nnz, 256));
    if (e == cudaSuccess) e = dafs::dev_malloc(&d_val, std::max<size_t>(sizeof(float) * nnz, 256));
    if (e == cudaSuccess) e = dafs::dev_malloc(&d_nz, sizeof(int64_t) * (units + 1));
    if (e == cudaSuccess) e = cudaMemcpy(d_nz, in->nz_off, sizeof(int64_t) * (units + 1), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && re) e = cudaMemcpy(d_rp, in->row_ptr, sizeof(int32_t) * re, cudaMemcpyDefault);
    if (e == cudaSuccess && nnz) e = cudaMemcpy(d_col, in->col, sizeof(uint32_t) * nnz, cudaMemcpyDefault);
    if (e == cudaSuccess && nnz) e = cudaMemcpy(d_val, in->val, sizeof(float) * nnz, cudaMemcpyDefault);
    if (e != cudaSuccess) return bail(fail(e == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM : DAFS_CUDA_ECUDA, "copy of the batch failed: %s", cudaGetErrorString(e)));
    const int rc = dafs::make_device_batch(units, nnz, in->row_off, d_nz, d_rp, d_col, d_val, nullptr, out);
    if (rc) return bail(rc);
    dafs::dev_free(d_nz);
    return DAFS_CUDA_OK;
  }
  HostBatch h;
  if (!h.alloc(units, re, nnz)) return fail(DAFS_CUDA_ENOMEM, "pinned allocation failed");
  std::memcpy(h.row_off, in->row_off, h.b_off);
  std::memcpy(h.nz_off, in->nz_off, h.b_off);
  cudaError_t e = re ? cudaMemcpy(h.row_ptr, in->row_ptr, h.b_row_ptr, cudaMemcpyDefault) : cudaSuccess;   // the input may be device-resident
  if (e == cudaSuccess && nnz) e = cudaMemcpy(h.col, in->col, h.b_col, cudaMemcpyDefault);
  if (e == cudaSuccess && nnz) e = cudaMemcpy(h.val, in->val, h.b_val, cudaMemcpyDefault);
  if (e != cudaSuccess) return fail(DAFS_CUDA_ECUDA, "copy of the batch failed: %s", cudaGetErrorString(e));
  h.give(out, units, nnz);
  return DAFS_CUDA_OK;
}

int grid_for(int64_t warps_wanted, int warps_per_block) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t blocks = (warps_wanted + warps_per_block - 1) / warps_per_block;
  return (int)std::max<int64_t>(1, std::min<int64_t>(blocks, (int64_t)sms * 16));
}

// device CSR of a relaxed batch
struct DevCsr { int64_t nnz = 0; int64_t* nz_off = nullptr; int* row_ptr = nullptr; uint32_t* col = nullptr; float* val = nullptr; };

}  // namespace

namespace dafs {

// pool -> CSR for `units` units whose output rows are out_row_off (device), then D2H
static int pct_finish(DevBufs& B, cudaStream_t st, int64_t units, int64_t rows, const int64_t* d_out_row_off,
                      const PctOut& O, int64_t nnz, const int64_t* h_row_off, dafs_csr_batch* out) {
  DevCsr C;
  int* d_unit_nnz = nullptr;
  DAFS_CUDA_TRY(B.get(&d_unit_nnz, units));
  DAFS_CUDA_TRY(B.get(&C.nz_off, units + 1));
  DAFS_CUDA_TRY(B.get(&C.row_ptr, rows + units));
  DAFS_CUDA_TRY(B.get(&C.col, nnz));
  DAFS_CUDA_TRY(B.get(&C.val, nnz));
  pct_unit_nnz_kernel<<<grid_for(units, 4), 128, 0, st>>>(units, d_out_row_off, O.row_cnt, d_unit_nnz);
  pct_scan_kernel<<<1, 1024, 0, st>>>(d_unit_nnz, units, C.nz_off);
  pct_gather_kernel<<<grid_for(units, 4), 128, 0, st>>>(units, d_out_row_off, O, C.nz_off, C.row_ptr, C.col, C.val);
  DAFS_CUDA_TRY(cudaGetLastError());
  if (g_device_out) {
    // the three arrays leave the call's buffer list and become the batch's
    const int rc = make_device_batch(units, nnz, h_row_off, C.nz_off, C.row_ptr, C.col, C.val, st, out);
    if (rc) return rc;
    B.release(C.row_ptr); B.release(C.col); B.release(C.val);
    return DAFS_CUDA_OK;
  }
  HostBatch h;
  if (!h.alloc(units, rows + units, nnz)) return fail(DAFS_CUDA_ENOMEM, "pinned allocation failed");
  std::memcpy(h.row_off, h_row_off, h.b_off);
  DAFS_CUDA_TRY(cudaMemcpyAsync(h.nz_off, C.nz_off, h.b_off, cudaMemcpyDeviceToHost, st));
  DAFS_CUDA_TRY(cudaMemcpyAsync(h.row_ptr, C.row_ptr, h.b_row_ptr, cudaMemcpyDeviceToHost, st));
  if (nnz) {
    DAFS_CUDA_TRY(cudaMemcpyAsync(h.col, C.col, h.b_col, cudaMemcpyDeviceToHost, st));
    DAFS_CUDA_TRY(cudaMemcpyAsync(h.val, C.val, h.b_val, cudaMemcpyDeviceToHost, st));
  }
  DAFS_CUDA_TRY(cudaStreamSynchronize(st));
  h.give(out, units, nnz);
  return DAFS_CUDA_OK;
}

}  // namespace dafs

using namespace dafs;

namespace {

struct StreamGuard {
  cudaStream_t s = nullptr;
  cudaEvent_t ev[6] = {};
  ~StreamGuard() {
    for (auto e : ev) if (e) cudaEventDestroy(e);
    if (s) cudaStreamDestroy(s);
  }
};

// pool for the kept entries of a relaxation; grows to the exact need on overflow
struct Pool {
  PctOut O{};
  int alloc(DevBufs& B, int64_t rows, int64_t cap) {
    DAFS_CUDA_TRY(B.get(&O.row_pool, rows));
    DAFS_CUDA_TRY(B.get(&O.row_cnt, rows));
    DAFS_CUDA_TRY(B.get(&O.pool_count, 1));
    return grow(B, cap);
  }
  int grow(DevBufs& B, int64_t cap) {
    if (O.pool_col) { B.drop(O.pool_col); B.drop(O.pool_val); O.pool_col = nullptr; O.pool_val = nullptr; }
    DAFS_CUDA_TRY(B.get(&O.pool_col, cap));
    DAFS_CUDA_TRY(B.get(&O.pool_val, cap));
    O.pool_cap = cap;
    return DAFS_CUDA_OK;
  }
};

}  // namespace

extern "C" int dafs_cuda_pct(int nseq, const int* len, const dafs_csr_batch* mp, const dafs_csr_batch* bp,
                             float w_pct_a, float w_pct_s, float* sim, dafs_csr_batch* mp_out,
                             dafs_csr_batch* bp_out) {
  if (mp_out) *mp_out = dafs_csr_batch{};
  if (bp_out) *bp_out = dafs_csr_batch{};
  for (float& v : g_ms) v = 0.0f;
  if (nseq < 1 || !len || !mp || !bp || !mp_out || !bp_out) return fail(DAFS_CUDA_EINVAL, "NULL argument or nseq < 1");
  const int n = nseq;
  const int64_t npairs = (int64_t)n * (n - 1) / 2;
  if (mp->n_units != npairs) return fail(DAFS_CUDA_EINVAL, "mp holds %lld units, %lld pairs expected", (long long)mp->n_units, (long long)npairs);
  if (bp->n_units != n) return fail(DAFS_CUDA_EINVAL, "bp holds %lld units, %d sequences expected", (long long)bp->n_units, n);
  int maxL = 0;
  for (int i = 0; i < n; ++i) {
    if (len[i] < 1 || len[i] > 65535) return fail(DAFS_CUDA_EINVAL, "len[%d] = %d (1 .. 65535)", i, len[i]);
    maxL = std::max(maxL, len[i]);
  }
  // geometry on the host: pair list, row-block offsets of U / T / outputs
  std::vector<int2> pairs((size_t)npairs);
  std::vector<int> ncols_mp((size_t)npairs);
  std::vector<int64_t> u_row_off(npairs + 1), t_row_off(npairs + 1), orow_mp(npairs + 1), b_row_off(n + 1), orow_bp(n + 1);
  {
    int64_t u = 0, ru = 0, rt = 0, ro = 0;
    for (int a = 0; a < n; ++a)
      for (int b = a + 1; b < n; ++b, ++u) {
        pairs[u] = make_int2(a, b); ncols_mp[u] = len[b];
        u_row_off[u] = ru; t_row_off[u] = rt; orow_mp[u] = ro;
        ru += len[a] + 1; rt += len[b] + 1; ro += len[a];
      }
    u_row_off[npairs] = ru; t_row_off[npairs] = rt; orow_mp[npairs] = ro;
    int64_t rb = 0, rob = 0;
    for (int a = 0; a < n; ++a) { b_row_off[a] = rb; orow_bp[a] = rob; rb += len[a] + 1; rob += len[a]; }
    b_row_off[n] = rb; orow_bp[n] = rob;
  }
  for (int64_t u = 0; u <= npairs; ++u)
    if (mp->row_off[u] != u_row_off[u]) return fail(DAFS_CUDA_EINVAL, "mp->row_off[%lld] does not match the sequence lengths", (long long)u);
  for (int a = 0; a <= n; ++a)
    if (bp->row_off[a] != b_row_off[a]) return fail(DAFS_CUDA_EINVAL, "bp->row_off[%d] does not match the sequence lengths", a);
  const int64_t mp_nnz = npairs ? mp->nz_off[npairs] : 0, bp_nnz = bp->nz_off[n];
  if (mp_nnz < 0 || bp_nnz < 0 || mp_nnz != mp->nnz || bp_nnz != bp->nnz) return fail(DAFS_CUDA_EINVAL, "nnz does not match nz_off");

  // the outputs are handed over only by a successful return: any error path frees what was filled so far
  struct OutGuard {
    dafs_csr_batch *a, *b; bool ok = false;
    ~OutGuard() { if (!ok) { dafs_cuda_free_batch(a); dafs_cuda_free_batch(b); } }
  } outs{mp_out, bp_out};
  StreamGuard G;
  DAFS_CUDA_TRY(cudaStreamCreateWithFlags(&G.s, cudaStreamNonBlocking));
  for (auto& e : G.ev) DAFS_CUDA_TRY(cudaEventCreate(&e));
  cudaStream_t st = G.s;
  DevBufs B;
  B.stream = st;
  PctArgs A{};
  A.n = n;
  int* d_len = nullptr; int2* d_pairs = nullptr; int* d_ncols_mp = nullptr; int* d_bad = nullptr; int* d_cursor = nullptr;
  int64_t *d_u_row_off = nullptr, *d_t_row_off = nullptr, *d_nz_off = nullptr, *d_b_row_off = nullptr, *d_b_nz_off = nullptr;
  int64_t *d_orow_mp = nullptr, *d_orow_bp = nullptr;
  int* d_u_row_ptr = nullptr; uint32_t* d_u_col = nullptr; float* d_u_val = nullptr;
  int* d_b_row_ptr = nullptr; uint32_t* d_b_col = nullptr; float* d_b_val = nullptr;
  DAFS_CUDA_TRY(B.get(&d_len, n));
  DAFS_CUDA_TRY(B.get(&d_pairs, npairs));
  DAFS_CUDA_TRY(B.get(&d_ncols_mp, npairs));
  DAFS_CUDA_TRY(B.get(&d_bad, 1));
  DAFS_CUDA_TRY(B.get(&d_u_row_off, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_t_row_off, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_nz_off, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_orow_mp, npairs + 1));
  DAFS_CUDA_TRY(B.get(&d_b_row_off, n + 1));
  DAFS_CUDA_TRY(B.get(&d_b_nz_off, n + 1));
  DAFS_CUDA_TRY(B.get(&d_orow_bp, n + 1));
  // U and T share one set of arrays: [U | T]
  const int64_t RU = u_row_off[npairs], RT = t_row_off[npairs];
  const int64_t RI = b_row_off[n], NI = b_row_off[n] - n;    // identity blocks: rows + 1 each, len entries
  DAFS_CUDA_TRY(B.get(&d_u_row_ptr, RU + RT + RI));
  DAFS_CUDA_TRY(B.get(&d_u_col, 2 * mp_nnz + NI));
  DAFS_CUDA_TRY(B.get(&d_u_val, 2 * mp_nnz + NI));
  A.t_row_ptr = d_u_row_ptr + RU; A.t_col = d_u_col + mp_nnz; A.t_val = d_u_val + mp_nnz;
  A.rp_all = d_u_row_ptr; A.col_all = d_u_col; A.val_all = d_u_val; A.ru = RU; A.nnz = mp_nnz;
  DAFS_CUDA_TRY(B.get(&d_cursor, RT));
  DAFS_CUDA_TRY(B.get(&A.desc, (size_t)n * n));
  DAFS_CUDA_TRY(B.get(&A.descT, (size_t)n * n));
  DAFS_CUDA_TRY(B.get(&A.head, RU + RT + b_row_off[n]));
  DAFS_CUDA_TRY(B.get(&A.ent_all, 2 * mp_nnz + NI));
  A.rt = RT;
  DAFS_CUDA_TRY(B.get(&d_b_row_ptr, b_row_off[n]));
  DAFS_CUDA_TRY(B.get(&d_b_col, bp_nnz));
  DAFS_CUDA_TRY(B.get(&d_b_val, bp_nnz));
  DAFS_CUDA_TRY(B.get(&A.sim, (size_t)n * n));

  DAFS_CUDA_TRY(cudaEventRecord(G.ev[0], st));
  auto h2d = [&](void* d, const void* h, size_t bytes) {
    return bytes ? cudaMemcpyAsync(d, h, bytes, cudaMemcpyDefault, st) : cudaSuccess;   // inputs may be device-resident batches
  };
  DAFS_CUDA_TRY(h2d(d_len, len, sizeof(int) * n));
  DAFS_CUDA_TRY(h2d(d_pairs, pairs.data(), sizeof(int2) * npairs));
  DAFS_CUDA_TRY(h2d(d_ncols_mp, ncols_mp.data(), sizeof(int) * npairs));
  DAFS_CUDA_TRY(h2d(d_u_row_off, u_row_off.data(), sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_t_row_off, t_row_off.data(), sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_orow_mp, orow_mp.data(), sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_nz_off, mp->nz_off, sizeof(int64_t) * (npairs + 1)));
  DAFS_CUDA_TRY(h2d(d_b_row_off, b_row_off.data(), sizeof(int64_t) * (n + 1)));
  DAFS_CUDA_TRY(h2d(d_orow_bp, orow_bp.data(), sizeof(int64_t) * (n + 1)));
  DAFS_CUDA_TRY(h2d(d_b_nz_off, bp->nz_off, sizeof(int64_t) * (n + 1)));
  DAFS_CUDA_TRY(h2d(d_u_row_ptr, mp->row_ptr, sizeof(int) * u_row_off[npairs]));
  DAFS_CUDA_TRY(h2d(d_u_col, mp->col, sizeof(uint32_t) * mp_nnz));
  DAFS_CUDA_TRY(h2d(d_u_val, mp->val, sizeof(float) * mp_nnz));
  DAFS_CUDA_TRY(h2d(d_b_row_ptr, bp->row_ptr, sizeof(int) * b_row_off[n]));
  DAFS_CUDA_TRY(h2d(d_b_col, bp->col, sizeof(uint32_t) * bp_nnz));
  DAFS_CUDA_TRY(h2d(d_b_val, bp->val, sizeof(float) * bp_nnz));
  DAFS_CUDA_TRY(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
  A.len = d_len;
  A.u_row_off = d_u_row_off; A.t_row_off = d_t_row_off; A.nz_off = d_nz_off;
  A.u_row_ptr = d_u_row_ptr; A.u_col = d_u_col; A.u_val = d_u_val;
  A.b_row_off = d_b_row_off; A.b_nz_off = d_b_nz_off; A.b_row_ptr = d_b_row_ptr; A.b_col = d_b_col; A.b_val = d_b_val;

  // inputs are indices into device arrays: check them before anything follows them
  if (npairs) pct_validate_kernel<<<grid_for(npairs, 4), 128, 0, st>>>(npairs, d_u_row_off, d_nz_off, d_u_row_ptr, d_u_col, d_ncols_mp, d_bad);
  pct_validate_kernel<<<grid_for(n, 4), 128, 0, st>>>(n, d_b_row_off, d_b_nz_off, d_b_row_ptr, d_b_col, d_len, d_bad);
  int bad = 0;
  DAFS_CUDA_TRY(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
  DAFS_CUDA_TRY(cudaStreamSynchronize(st));
  if (bad) return fail(DAFS_CUDA_EINVAL, "mp / bp is not a well-formed row-sparse batch (row_ptr, ascending columns, column range)");

  // K6, K7
  DAFS_CUDA_TRY(cudaEventRecord(G.ev[1], st));
  pct_desc_kernel<<<(n * n + 255) / 256, 256, 0, st>>>(A);
  pct_identity_kernel<<<n, 128, 0, st>>>(A, d_u_row_ptr, d_u_col, d_u_val);
  if (npairs) pct_transpose_kernel<<<grid_for(npairs, 4), 128, 0, st>>>(A, npairs, d_pairs, d_cursor);
  DAFS_CUDA_TRY(cudaEventRecord(G.ev[2], st));
  {
    const int W = maxL + 2;
    int grid = grid_for(npairs, 4);
    while (grid > 1 && (size_t)grid * 4 * 5 * W * sizeof(float) > (size_t(256) << 20)) grid /= 2;
    float* d_scr = nullptr;
    DAFS_CUDA_TRY(B.get(&d_scr, (size_t)grid * 4 * 5 * W));
    if (npairs) pct_similarity_kernel<<<grid, 128, 0, st>>>(A, npairs, d_pairs, d_scr, W);
    pct_sim_diag_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, A.sim);
  }
  DAFS_CUDA_TRY(cudaEventRecord(G.ev[3], st));
  DAFS_CUDA_TRY(cudaGetLastError());
  if (sim) DAFS_CUDA_TRY(cudaMemcpyAsync(sim, A.sim, sizeof(float) * (size_t)n * n, cudaMemcpyDeviceToHost, st));

  // accumulator rows in shared memory: warps per block by the longest sequence
  const int accW = (maxL + 31) & ~31;
  int wpb = 4;
  while (wpb > 1 && (size_t)wpb * accW * sizeof(float) > size_t(200) << 10) wpb /= 2;
  const size_t smem = (size_t)wpb * accW * sizeof(float);
  if (smem > (size_t(227) << 10)) return fail(DAFS_CUDA_EINVAL, "sequence of length %d is too long for the consistency kernels", maxL);
  if (smem > (48 << 10)) {
    DAFS_CUDA_TRY(cudaFuncSetAttribute(pct_relax_bp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }

  // K8
  int rc = DAFS_CUDA_OK;
  if (w_pct_s != 0.0f) {
    float *d_w = nullptr, *d_sw = nullptr;
    DAFS_CUDA_TRY(B.get(&d_w, (size_t)n * n));
    DAFS_CUDA_TRY(B.get(&d_sw, n));
    pct_weights_bp_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, A.sim, w_pct_s, d_w, d_sw);
    Pool P;
    const int64_t rows = orow_bp[n];
    if ((rc = P.alloc(B, rows, std::max<int64_t>(2 * bp_nnz, 1024)))) return rc;
    int64_t need = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
      DAFS_CUDA_TRY(cudaMemsetAsync(P.O.pool_count, 0, sizeof(unsigned long long), st));
      pct_relax_bp_kernel<<<grid_for(rows, wpb), 32 * wpb, smem, st>>>(A, d_orow_bp, d_w, d_sw, P.O, accW);
      DAFS_CUDA_TRY(cudaGetLastError());
      DAFS_CUDA_TRY(cudaMemcpyAsync(&need, P.O.pool_count, sizeof need, cudaMemcpyDeviceToHost, st));
      DAFS_CUDA_TRY(cudaStreamSynchronize(st));
      if (need <= P.O.pool_cap) break;
      if (attempt == 1) return fail(DAFS_CUDA_ECUDA, "relax_bp pool overflow after regrow");
      if ((rc = P.grow(B, need))) return rc;
    }
    DAFS_CUDA_TRY(cudaEventRecord(G.ev[4], st));
    if ((rc = pct_finish(B, st, n, rows, d_orow_bp, P.O, need, b_row_off.data(), bp_out))) return rc;
  } else {
    DAFS_CUDA_TRY(cudaEventRecord(G.ev[4], st));
    if ((rc = copy_batch(bp, bp_out))) return rc;
  }

  // K9
  cudaEvent_t ev_mp0 = nullptr, ev_mp1 = nullptr;
  struct EvGuard { cudaEvent_t& a; cudaEvent_t& b; ~EvGuard() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); } } eg{ev_mp0, ev_mp1};
  DAFS_CUDA_TRY(cudaEventCreate(&ev_mp0));
  DAFS_CUDA_TRY(cudaEventCreate(&ev_mp1));
  DAFS_CUDA_TRY(cudaEventRecord(ev_mp0, st));
  if (w_pct_a != 0.0f && npairs) {
    float *d_w = nullptr, *d_sw = nullptr;
    DAFS_CUDA_TRY(B.get(&d_w, (size_t)npairs * n));
    DAFS_CUDA_TRY(B.get(&d_sw, npairs));
    pct_weights_mp_kernel<<<(int)((npairs + 127) / 128), 128, 0, st>>>(n, npairs, d_pairs, A.sim, w_pct_a, d_w, d_sw);
    // tasks = (pair, block of 32 rows) in tiles of 16 x 16 pairs: a tile reads mp_[x][.] of
    // 16 x's and mp_[.][y] of 16 y's, which stay in L2 while its tasks run
    std::vector<int2> tasks;
    tasks.reserve((size_t)npairs * ((maxL + 31) / 32));
    int TILE = (double)mp_nnz > 2.5 * (double)orow_mp[npairs] ? 8 : 16;
    if (const char* t = getenv("DAFS_PCT_TILE")) TILE = std::max(1, atoi(t));
    for (int x0 = 0; x0 < n; x0 += TILE)
      for (int y0 = x0; y0 < n; y0 += TILE)
        for (int x = x0; x < std::min(n, x0 + TILE); ++x)
          for (int rb = 0; rb < (len[x] + 31) / 32; ++rb)     // the warps of a CTA share (x, rows)
            for (int y = std::max(y0, x + 1); y < std::min(n, y0 + TILE); ++y)
              tasks.push_back(make_int2((int)((int64_t)x * n - (int64_t)x * (x + 1) / 2 + (y - x - 1)), rb));
    const int64_t ntasks = (int64_t)tasks.size();
    int2* d_tasks = nullptr;
    DAFS_CUDA_TRY(B.get(&d_tasks, ntasks));
    DAFS_CUDA_TRY(h2d(d_tasks, tasks.data(), sizeof(int2) * ntasks));
    // Two kernels: rows of one or two entries (related sequences) go through the row heads, a
    // thread per row; rows with many entries (unrelated sequences) take G = 8 lanes per row with
    // the accumulator rows in shared memory.
    bool dense = (double)mp_nnz > 2.5 * (double)orow_mp[npairs];
    if (const char* k = getenv("DAFS_PCT_MP_KERNEL")) dense = k[0] == 'd';
    int G = 8;                                               // lanes per row
    if (const char* k = getenv("DAFS_PCT_G")) G = atoi(k) == 4 ? 4 : 8;
    int wave = 0;                                            // 1: one launch per wave of resident warps
    if (const char* k = getenv("DAFS_PCT_WAVE")) wave = atoi(k);
    const int RW = 32 / G, UPT = 32 / RW;                    // rows per warp, units per task of 32 rows
    const size_t dense_smem_want = (size_t)8 * RW * accW * sizeof(float);
    bool dense_global = dense_smem_want > (size_t(200) << 10);
    if (const char* k = getenv("DAFS_PCT_MP_KERNEL")) if (k[0] == 'd' && k[1] == 'g') dense_global = true;   // "dg": tests
    if (dense_global) wave = 1;                              // the global scratch is sized for one wave
    const size_t dense_smem = dense_global ? 0 : dense_smem_want;
    int mp_grid = 0;
    size_t acc_floats_per_warp = dense ? (dense_global ? (size_t)RW * accW : 0) : (size_t)maxL * 32;
    {
      int dev = 0, sms = 148, per_sm = 1;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (dense) {
        if (dense_smem > (48 << 10)) {
          DAFS_CUDA_TRY(cudaFuncSetAttribute(pct_relax_mp_dense_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dense_smem));
          DAFS_CUDA_TRY(cudaFuncSetAttribute(pct_relax_mp_dense_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dense_smem));
        }
        if (G == 4) {
          if (dense_global) DAFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pct_relax_mp_dense_kernel<4, false>, 256, 0));
          else DAFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pct_relax_mp_dense_kernel<4, true>, 256, dense_smem));
        } else {
          if (dense_global) DAFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pct_relax_mp_dense_kernel<8, false>, 256, 0));
          else DAFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pct_relax_mp_dense_kernel<8, true>, 256, dense_smem));
        }
      } else {
        DAFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pct_relax_mp_kernel, 256, 0));
      }
      mp_grid = (int)std::max<int64_t>(1, std::min<int64_t>((ntasks + 7) / 8, (int64_t)sms * std::max(per_sm, 1)));
    }
    if (const char* g = getenv("DAFS_PCT_MP_GRID")) mp_grid = std::max(1, atoi(g));
    while (mp_grid > 1 && (size_t)mp_grid * 8 * acc_floats_per_warp * sizeof(float) > (size_t(2) << 30)) mp_grid /= 2;
    const size_t acc_warps = (size_t)mp_grid * 8;
    float* d_acc = nullptr;
    DAFS_CUDA_TRY(B.get(&d_acc, acc_warps * acc_floats_per_warp));
    Pool P;
    const int64_t rows = orow_mp[npairs];
    if ((rc = P.alloc(B, rows, std::max<int64_t>(2 * mp_nnz, 1024)))) return rc;
    int64_t need = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
      cudaError_t e = cudaMemsetAsync(P.O.pool_count, 0, sizeof(unsigned long long), st);
      if (e == cudaSuccess) {
        if (dense) {
          // one launch over all units, or one per wave of resident warps (global scratch)
          const int64_t units = ntasks * UPT, per_wave = wave ? (int64_t)mp_grid * 8 : units;
          for (int64_t u0 = 0; u0 < units; u0 += per_wave) {
            const int gridw = (int)((std::min(per_wave, units - u0) + 7) / 8);
            if (G == 4) {
              if (dense_global) pct_relax_mp_dense_kernel<4, false><<<gridw, 256, 0, st>>>(A, u0, ntasks, d_tasks, d_pairs, d_orow_mp, d_w, d_sw, P.O, d_acc, accW);
              else pct_relax_mp_dense_kernel<4, true><<<gridw, 256, dense_smem, st>>>(A, u0, ntasks, d_tasks, d_pairs, d_orow_mp, d_w, d_sw, P.O, nullptr, accW);
            } else {
              if (dense_global) pct_relax_mp_dense_kernel<8, false><<<gridw, 256, 0, st>>>(A, u0, ntasks, d_tasks, d_pairs, d_orow_mp, d_w, d_sw, P.O, d_acc, accW);
              else pct_relax_mp_dense_kernel<8, true><<<gridw, 256, dense_smem, st>>>(A, u0, ntasks, d_tasks, d_pairs, d_orow_mp, d_w, d_sw, P.O, nullptr, accW);
            }
          }
        }
        else
          pct_relax_mp_kernel<<<mp_grid, 256, 0, st>>>(A, ntasks, d_tasks, d_pairs, d_orow_mp, d_w, d_sw, P.O, d_acc, maxL);
        e = cudaGetLastError();
      }
      if (e == cudaSuccess) e = cudaMemcpyAsync(&need, P.O.pool_count, sizeof need, cudaMemcpyDeviceToHost, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) { return fail(DAFS_CUDA_ECUDA, "relax_mp: %s", cudaGetErrorString(e)); }
      if (need <= P.O.pool_cap) break;
      if (attempt == 1) { return fail(DAFS_CUDA_ECUDA, "relax_mp pool overflow after regrow"); }
      if ((rc = P.grow(B, need))) return rc;
    }
    cudaEventRecord(ev_mp1, st);
    if ((rc = pct_finish(B, st, npairs, rows, d_orow_mp, P.O, need, u_row_off.data(), mp_out))) return rc;
  } else {
    cudaEventRecord(ev_mp1, st);
    if ((rc = copy_batch(mp, mp_out))) return rc;
  }
  cudaEventRecord(G.ev[5], st);
  if (cudaStreamSynchronize(st) != cudaSuccess) {
    return fail(DAFS_CUDA_ECUDA, "consistency transformation failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  cudaEventElapsedTime(&g_ms[0], G.ev[1], G.ev[2]);
  cudaEventElapsedTime(&g_ms[1], G.ev[2], G.ev[3]);
  cudaEventElapsedTime(&g_ms[2], G.ev[3], G.ev[4]);
  cudaEventElapsedTime(&g_ms[3], ev_mp0, ev_mp1);
  cudaEventElapsedTime(&g_ms[4], G.ev[0], G.ev[5]);
  outs.ok = true;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_pct_device(int nseq, const int* len, const dafs_csr_batch* mp, const dafs_csr_batch* bp,
                                    float w_pct_a, float w_pct_s, float* sim, dafs_csr_batch* mp_out,
                                    dafs_csr_batch* bp_out) {
  g_device_out = true;
  const int rc = dafs_cuda_pct(nseq, len, mp, bp, w_pct_a, w_pct_s, sim, mp_out, bp_out);
  g_device_out = false;
  return rc;
}

extern "C" void dafs_cuda_pct_last_ms(float out5[5]) {
  for (int k = 0; k < 5; ++k) out5[k] = g_ms[k];
}
