// See dafs_cuda_models.h.  Everything here is host glue between the reference's
// std::vector types (src/typedefs.h:25-43) and the plain-C ABI of libdafs_cuda.
#include "dafs_cuda_models.h"

#include <algorithm>

#include "dafs_cuda.h"

namespace {

[[noreturn]] void fail(const char* what) {
  throw std::string(what) + ": " + dafs_cuda_last_error();
}

// One unit of a dafs_csr_batch -> MP / BP rows (vector<vector<pair<uint,float>>>).
void unpack_unit(const dafs_csr_batch& b, int64_t u, std::vector<SV>& rows) {
  const int64_t r0 = b.row_off[u], nrows = b.row_off[u + 1] - r0 - 1;
  const int32_t* rp = b.row_ptr + r0;
  const uint32_t* col = b.col + b.nz_off[u];
  const float* val = b.val + b.nz_off[u];
  rows.assign(nrows, SV());
  for (int64_t r = 0; r < nrows; ++r) {
    rows[r].reserve(rp[r + 1] - rp[r]);
    for (int32_t k = rp[r]; k < rp[r + 1]; ++k) rows[r].push_back(std::make_pair(col[k], val[k]));
  }
}

std::vector<float> flatten(const VVF& m, size_t cols) {
  std::vector<float> f(m.size() * cols);
  for (size_t i = 0; i < m.size(); ++i) std::copy(m[i].begin(), m[i].end(), f.begin() + i * cols);
  return f;
}

// MP / BP rows of several units -> the arrays of a dafs_csr_batch (caller-owned vectors).
struct PackedBatch {
  std::vector<int64_t> row_off, nz_off;
  std::vector<int32_t> row_ptr;
  std::vector<uint32_t> col;
  std::vector<float> val;
  dafs_csr_batch view;
  void add(const std::vector<SV>& rows) {
    if (row_off.empty()) { row_off.push_back(0); nz_off.push_back(0); }
    int32_t k = 0;
    row_ptr.push_back(0);
    for (const SV& r : rows) {
      for (const auto& e : r) { col.push_back(e.first); val.push_back(e.second); }
      k += (int32_t)r.size();
      row_ptr.push_back(k);
    }
    row_off.push_back((int64_t)row_ptr.size());
    nz_off.push_back((int64_t)col.size());
  }
  const dafs_csr_batch* finish() {
    if (row_off.empty()) { row_off.push_back(0); nz_off.push_back(0); }
    view = dafs_csr_batch();
    view.n_units = (int64_t)row_off.size() - 1; view.nnz = (int64_t)col.size();
    view.row_off = row_off.data(); view.row_ptr = row_ptr.data(); view.nz_off = nz_off.data();
    view.col = col.data(); view.val = val.data();
    return &view;
  }
};

}  // namespace

// ---- Align::Model ---------------------------------------------------------------
CudaProbCons::CudaProbCons(float th, int ngpu) : Align::Model(th), ngpu_(ngpu) {
  if (dafs_cuda_init(ngpu) != DAFS_CUDA_OK) fail("dafs_cuda_init");
}

void CudaProbCons::calculate(const std::string& seq1, const std::string& seq2, MP& mp) {
  const char* seq[2] = {seq1.data(), seq2.data()};
  const int len[2] = {(int)seq1.size(), (int)seq2.size()};
  const int32_t pi = 0, pj = 1;
  dafs_csr_batch b;
  if (dafs_cuda_pairhmm_pairs(2, seq, len, 1, &pi, &pj, threshold(), &b) != DAFS_CUDA_OK) fail("dafs_cuda_pairhmm_pairs");
  unpack_unit(b, 0, mp);
  dafs_cuda_free_batch(&b);
}

void CudaProbCons::calculate(const std::vector<Fasta>& fa, std::vector<std::vector<MP> >& mp) {
  const uint N = fa.size();
  std::vector<const char*> seq(N);
  std::vector<int> len(N);
  for (uint i = 0; i != N; ++i) { seq[i] = fa[i].seq().data(); len[i] = (int)fa[i].size(); }
  dafs_csr_batch b;
  if (dafs_cuda_pairhmm_all((int)N, seq.data(), len.data(), threshold(), ngpu_, &b) != DAFS_CUDA_OK) fail("dafs_cuda_pairhmm_all");
  mp.resize(N, std::vector<MP>(N));
  int64_t u = 0;
  for (uint i = 0; i != N; ++i) {
    mp[i][i].resize(fa[i].size());
    for (uint x = 0; x != fa[i].size(); ++x) mp[i][i][x].push_back(std::make_pair(x, 1.0f));
    for (uint j = i + 1; j != N; ++j) unpack_unit(b, u++, mp[i][j]);
  }
  dafs_cuda_free_batch(&b);
}

// ---- Fold::Model -----------------------------------------------------------------
CudaBoltzmann::CudaBoltzmann(float th, int ngpu) : Fold::Model(th), ngpu_(ngpu) {
  if (dafs_cuda_init(ngpu) != DAFS_CUDA_OK) fail("dafs_cuda_init");
}

void CudaBoltzmann::calculate(const std::string& seq, BP& bp) {
  const char* s[1] = {seq.data()};
  const int len[1] = {(int)seq.size()};
  dafs_csr_batch b;
  if (dafs_cuda_mccaskill_all(1, s, len, threshold(), ngpu_, &b) != DAFS_CUDA_OK) fail("dafs_cuda_mccaskill_all");
  unpack_unit(b, 0, bp);
  dafs_cuda_free_batch(&b);
}

// RNAfold::calculate(seq, str, bp) (src/fold.cpp:118-166): the constraint goes to the device in the
// reference's own alphabet ('?' free, '.' unpaired, '(' ')' a given pair)
void CudaBoltzmann::calculate(const std::string& seq, const std::string& str, BP& bp) {
  if (seq.size() != str.size()) throw std::string("constraint and sequence differ in length");
  const char* s[1] = {seq.data()};
  const char* c[1] = {str.data()};
  const int len[1] = {(int)seq.size()};
  dafs_csr_batch b;
  if (dafs_cuda_mccaskill_constrained(1, s, len, c, threshold(), &b) != DAFS_CUDA_OK) fail("dafs_cuda_mccaskill_constrained");
  unpack_unit(b, 0, bp);
  dafs_cuda_free_batch(&b);
}

void CudaBoltzmann::calculate(const std::vector<Fasta>& fa, std::vector<BP>& bp) {
  const uint N = fa.size();
  std::vector<const char*> seq(N);
  std::vector<int> len(N);
  for (uint i = 0; i != N; ++i) { seq[i] = fa[i].seq().data(); len[i] = (int)fa[i].size(); }
  dafs_csr_batch b;
  if (dafs_cuda_mccaskill_all((int)N, seq.data(), len.data(), threshold(), ngpu_, &b) != DAFS_CUDA_OK) fail("dafs_cuda_mccaskill_all");
  bp.resize(N);
  for (uint i = 0; i != N; ++i) unpack_unit(b, i, bp[i]);
  dafs_cuda_free_batch(&b);
}

// ---- decoders ----------------------------------------------------------------------
float CudaNussinov::decode(float w, const VVF& p, const VVF& q, VU& ss) {
  const int L = (int)p.size();
  std::vector<float> fp = flatten(p, L), fq = flatten(q, L);
  ss.resize(L);
  float s = 0;
  if (dafs_cuda_nussinov_decode(L, th_, w, fp.data(), fq.data(), ss.data(), &s) != DAFS_CUDA_OK) fail("dafs_cuda_nussinov_decode");
  return s;
}

float CudaNussinov::decode(const VVF& p, VU& ss, std::string& str) {
  const int L = (int)p.size();
  std::vector<float> fp = flatten(p, L);
  ss.resize(L);
  float s = 0;
  if (dafs_cuda_nussinov_decode(L, th_, 1.0f, fp.data(), nullptr, ss.data(), &s) != DAFS_CUDA_OK) fail("dafs_cuda_nussinov_decode");
  make_brackets(ss, str);
  return s;
}

void CudaNussinov::make_brackets(const VU& ss, std::string& str) const {   // src/nussinov.cpp:395-413
  str.assign(ss.size(), '.');
  for (uint i = 0; i != ss.size(); ++i)
    if (ss[i] != -1u) { str[i] = Fold::Decoder::left_brackets[0]; str[ss[i]] = Fold::Decoder::right_brackets[0]; }
}

float CudaNeedlemanWunsch::decode(const VVF& p, const VVF& q, VU& al) const {
  const int L1 = (int)p.size(), L2 = (int)p[0].size();
  std::vector<float> fp = flatten(p, L2), fq = flatten(q, L2);
  al.resize(L1);
  float s = 0;
  if (dafs_cuda_nw_decode(L1, L2, th_, fp.data(), fq.data(), al.data(), &s) != DAFS_CUDA_OK) fail("dafs_cuda_nw_decode");
  return s;
}

float CudaNeedlemanWunsch::decode(const VVF& p, VU& al) const {
  const int L1 = (int)p.size(), L2 = (int)p[0].size();
  std::vector<float> fp = flatten(p, L2);
  al.resize(L1);
  float s = 0;
  if (dafs_cuda_nw_decode(L1, L2, th_, fp.data(), nullptr, al.data(), &s) != DAFS_CUDA_OK) fail("dafs_cuda_nw_decode");
  return s;
}

// ---- dual decomposition ---------------------------------------------------------------
namespace dafs_cuda {

void solve_by_dd_batch(const DDOptions& o, const std::vector<DDInput>& in, std::vector<DDOutput>& out) {
  const size_t n = in.size();
  std::vector<std::vector<float> > fx(n), fy(n), fz(n);
  std::vector<dafs_dd_problem> prob(n);
  std::vector<dafs_dd_result> res(n);
  out.resize(n);
  const float min_th_s = *std::min_element(o.th_s.begin(), o.th_s.end());   // src/dafs.cpp:1018
  for (size_t k = 0; k < n; ++k) {
    const int L1 = (int)in[k].p_x->size(), L2 = (int)in[k].p_y->size();
    fx[k] = flatten(*in[k].p_x, L1); fy[k] = flatten(*in[k].p_y, L2); fz[k] = flatten(*in[k].p_z, L2);
    dafs_dd_problem& q = prob[k];
    q.L1 = L1; q.L2 = L2; q.N1 = (int)in[k].n1; q.N2 = (int)in[k].n2;
    q.p_x = fx[k].data(); q.p_y = fy[k].data(); q.p_z = fz[k].data();
    q.w = o.w; q.th_s = o.th_s[0]; q.min_th_s = min_th_s; q.th_a = o.th_a; q.eta0 = o.eta0; q.t_max = (int)o.t_max;
    out[k].x.assign(L1, -1u); out[k].y.assign(L2, -1u); out[k].z.assign(L1, -1u);
    res[k] = dafs_dd_result();
    res[k].x = out[k].x.data(); res[k].y = out[k].y.data(); res[k].z = out[k].z.data();
  }
  if (dafs_cuda_dd_solve_batch((int)n, prob.data(), res.data()) != DAFS_CUDA_OK) fail("dafs_cuda_dd_solve_batch");
  for (size_t k = 0; k < n; ++k) {
    if (res[k].status != 0) throw std::string("dafs_cuda_dd_solve_batch: alignment traceback left the envelope");
    out[k].outcome.score = res[k].score; out[k].outcome.iters = (unsigned)res[k].iters;
    out[k].outcome.violated = res[k].violated;
  }
}

DDOutcome solve_by_dd(const DDOptions& o, VU& x, VU& y, VU& z, const VVF& p_x, const VVF& p_y,
                      const VVF& p_z, const ALN& aln1, const ALN& aln2) {
  std::vector<DDInput> in(1);
  in[0].p_x = &p_x; in[0].p_y = &p_y; in[0].p_z = &p_z;
  in[0].n1 = (unsigned)aln1.size(); in[0].n2 = (unsigned)aln2.size();                // src/dafs.cpp:1013-1014
  std::vector<DDOutput> out;
  solve_by_dd_batch(o, in, out);
  x.swap(out[0].x); y.swap(out[0].y); z.swap(out[0].z);
  return out[0].outcome;
}

void consistency_transform(const std::vector<Fasta>& fa, std::vector<std::vector<MP> >& mp, std::vector<BP>& bp,
                           float w_pct_a, float w_pct_s, VVF& sim) {
  const uint N = fa.size();
  std::vector<int> len(N);
  PackedBatch pm, pb;
  for (uint i = 0; i != N; ++i) {
    len[i] = (int)fa[i].size();
    pb.add(bp[i]);
    for (uint j = i + 1; j != N; ++j) pm.add(mp[i][j]);
  }
  std::vector<float> s((size_t)N * N);
  dafs_csr_batch om, ob;
  if (dafs_cuda_pct((int)N, len.data(), pm.finish(), pb.finish(), w_pct_a, w_pct_s, s.data(), &om, &ob) != DAFS_CUDA_OK)
    fail("dafs_cuda_pct");
  sim.assign(N, VF(N));
  int64_t u = 0;
  for (uint i = 0; i != N; ++i) {
    std::copy(s.begin() + (size_t)i * N, s.begin() + (size_t)(i + 1) * N, sim[i].begin());
    unpack_unit(ob, i, bp[i]);
    for (uint j = i + 1; j != N; ++j, ++u) {
      unpack_unit(om, u, mp[i][j]);
      // transpose_mp (src/dafs.cpp:155-167): rows of mp[i][j] in order give ascending rows per column
      MP& t = mp[j][i];
      t.assign(fa[j].size(), SV());
      for (uint r = 0; r != mp[i][j].size(); ++r)
        for (const auto& e : mp[i][j][r]) t[e.first].push_back(std::make_pair(r, e.second));
    }
  }
  dafs_cuda_free_batch(&om);
  dafs_cuda_free_batch(&ob);
}

void build_tree(const VVF& sim, std::vector<TreeNode>& tree) {
  const uint N = sim.size();
  std::vector<float> s = flatten(sim, N);
  std::vector<int32_t> left(2 * N - 1), right(2 * N - 1);
  std::vector<float> score(2 * N - 1);
  if (dafs_cuda_guide_tree((int)N, s.data(), left.data(), right.data(), score.data()) != DAFS_CUDA_OK)
    fail("dafs_cuda_guide_tree");
  tree.resize(2 * N - 1);
  for (uint k = 0; k != 2 * N - 1; ++k)      // leaves keep (0.0, (-1u, -1u)) like src/dafs.cpp:452
    tree[k] = std::make_pair(score[k], std::make_pair((uint)left[k], (uint)right[k]));
}

float align_tree(const DDOptions& o, const std::vector<Fasta>& fa, const std::vector<std::vector<MP> >& mp,
                 const std::vector<BP>& bp, const std::vector<TreeNode>& tree, VU& ss, ALN& aln) {
  const uint N = fa.size();
  std::vector<int> len(N);
  PackedBatch pm, pb;
  for (uint i = 0; i != N; ++i) {
    len[i] = (int)fa[i].size();
    pb.add(bp[i]);
    for (uint j = i + 1; j != N; ++j) pm.add(mp[i][j]);
  }
  std::vector<int32_t> left(tree.size()), right(tree.size());
  for (size_t k = 0; k != tree.size(); ++k) { left[k] = (int32_t)tree[k].second.first; right[k] = (int32_t)tree[k].second.second; }
  dafs_prog_params prm;
  prm.w = o.w; prm.th_s = o.th_s[0]; prm.min_th_s = *std::min_element(o.th_s.begin(), o.th_s.end());
  prm.th_a = o.th_a; prm.eta0 = o.eta0; prm.t_max = (int32_t)o.t_max;
  dafs_prog_result r;
  if (dafs_cuda_progressive((int)N, len.data(), pm.finish(), pb.finish(), left.data(), right.data(), &prm, &r) != DAFS_CUDA_OK)
    fail("dafs_cuda_progressive");
  aln.resize(r.n_rows);
  for (int q = 0; q < r.n_rows; ++q) {
    aln[q].first = (uint)r.seq[q];
    aln[q].second.assign(r.bits + (size_t)q * r.L, r.bits + (size_t)(q + 1) * r.L);
  }
  if (r.ss) ss.assign(r.ss, r.ss + r.L);      // a single sequence leaves ss untouched, like align()
  const float s = r.score;
  dafs_cuda_free_prog_result(&r);
  return s;
}

// Everything between :1787 and :1838 of DAFS::run() in one call, with the posteriors resident in HBM
// from the models to the alignment: McCaskill and pair-HMM plans hand their results over as device
// batches, the consistency transformation reads and writes device batches, the progressive stage reads
// them.  Only sim (N x N floats), the tree and the alignment reach the host.
float align_from_sequences(const DDOptions& o, const std::vector<Fasta>& fa, float th_bp, float w_pct_a, float w_pct_s,
                           VVF& sim, std::vector<TreeNode>& tree, VU& ss, ALN& aln) {
  const uint N = fa.size();
  std::vector<const char*> seq(N);
  std::vector<int> len(N);
  for (uint i = 0; i != N; ++i) { seq[i] = fa[i].seq().data(); len[i] = (int)fa[i].size(); }
  std::vector<int32_t> pi, pj;
  for (uint i = 0; i != N; ++i) for (uint j = i + 1; j != N; ++j) { pi.push_back((int32_t)i); pj.push_back((int32_t)j); }
  struct Batches {
    dafs_csr_batch b[4] = {};
    dafs_mccaskill_plan* pm = nullptr; dafs_pairhmm_plan* pp = nullptr;
    ~Batches() { for (auto& x : b) dafs_cuda_free_batch(&x); dafs_cuda_mccaskill_plan_destroy(pm); dafs_cuda_pairhmm_plan_destroy(pp); }
  } B;
  if (dafs_cuda_mccaskill_plan_create((int)N, seq.data(), len.data(), th_bp, &B.pm) != DAFS_CUDA_OK) fail("dafs_cuda_mccaskill_plan_create");
  if (dafs_cuda_pairhmm_plan_create((int)N, seq.data(), len.data(), (int64_t)pi.size(), pi.data(), pj.data(), o.th_a, &B.pp) != DAFS_CUDA_OK)
    fail("dafs_cuda_pairhmm_plan_create");
  if (dafs_cuda_mccaskill_plan_run(B.pm) != DAFS_CUDA_OK || dafs_cuda_pairhmm_plan_run(B.pp) != DAFS_CUDA_OK) fail("plan_run");   // the two stages overlap on the device
  if (dafs_cuda_mccaskill_plan_fetch_device(B.pm, &B.b[0]) != DAFS_CUDA_OK) fail("dafs_cuda_mccaskill_plan_fetch_device");
  if (dafs_cuda_pairhmm_plan_fetch_device(B.pp, &B.b[1]) != DAFS_CUDA_OK) fail("dafs_cuda_pairhmm_plan_fetch_device");
  std::vector<float> s((size_t)N * N);
  if (dafs_cuda_pct_device((int)N, len.data(), &B.b[1], &B.b[0], w_pct_a, w_pct_s, s.data(), &B.b[2], &B.b[3]) != DAFS_CUDA_OK) fail("dafs_cuda_pct_device");
  sim.assign(N, VF(N));
  for (uint i = 0; i != N; ++i) std::copy(s.begin() + (size_t)i * N, s.begin() + (size_t)(i + 1) * N, sim[i].begin());
  build_tree(sim, tree);
  std::vector<int32_t> left(tree.size()), right(tree.size());
  for (size_t k = 0; k != tree.size(); ++k) { left[k] = (int32_t)tree[k].second.first; right[k] = (int32_t)tree[k].second.second; }
  dafs_prog_params prm;
  prm.w = o.w; prm.th_s = o.th_s[0]; prm.min_th_s = *std::min_element(o.th_s.begin(), o.th_s.end());
  prm.th_a = o.th_a; prm.eta0 = o.eta0; prm.t_max = (int32_t)o.t_max;
  dafs_prog_result r;
  if (dafs_cuda_progressive((int)N, len.data(), &B.b[2], &B.b[3], left.data(), right.data(), &prm, &r) != DAFS_CUDA_OK)
    fail("dafs_cuda_progressive");
  aln.resize(r.n_rows);
  for (int q = 0; q < r.n_rows; ++q) {
    aln[q].first = (uint)r.seq[q];
    aln[q].second.assign(r.bits + (size_t)q * r.L, r.bits + (size_t)(q + 1) * r.L);
  }
  if (r.ss) ss.assign(r.ss, r.ss + r.L);
  const float sc = r.score;
  dafs_cuda_free_prog_result(&r);
  return sc;
}

}  // namespace dafs_cuda
