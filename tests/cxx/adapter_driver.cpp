// TEST INFRASTRUCTURE.  Links the C++ adapters of dafs_b200/host against the reference's own
// classes (compiled unmodified into oracle/_ref/libdafs_ref.so) and libdafs_cuda.so, and checks on
// tests/golden/RF00005_0.fa that each adapter returns exactly what the CPU class it replaces
// returns:  CudaProbCons vs ProbCons (src/align.cpp:54-79), CudaNussinov vs SparseNussinov
// (src/nussinov.cpp:207-413), CudaNeedlemanWunsch vs SparseNeedlemanWunsch
// (src/needleman_wunsch.cpp:198-422), dafs_cuda::solve_by_dd vs DAFS::solve_by_dd
// (src/dafs.cpp:1006-1295, lifted at build time into the same library).  Built by
// oracle/Makefile into oracle/_ref/adapter_driver (this container only: it needs the reference
// headers); run on the GPU box by tests/test_host_adapters_gpu.py.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "align.h"
#include "dafs_cuda_models.h"
#include "fold.h"
#include "needleman_wunsch.h"
#include "nussinov.h"

// src/fold.cpp:60-68 (fold.cpp itself needs ViennaRNA and is not part of the oracle library)
void Fold::Model::calculate(const std::vector<Fasta>& fa, std::vector<BP>& bp) {
  bp.resize(fa.size());
  for (uint i = 0; i != fa.size(); ++i) this->calculate(fa[i].seq(), bp[i]);
}

extern "C" float ref_dd_solve(int L1, int L2, int N1, int N2, const float* p_x, const float* p_y, const float* p_z,
                              float w, float th_s, float th_a, float eta0, int t_max, unsigned* x, unsigned* y,
                              unsigned* z, int* iters, unsigned* violated);

static int g_fail = 0;
#define CHECK(cond, ...) do { if (!(cond)) { ++g_fail; std::fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); std::fprintf(stderr, __VA_ARGS__); std::fprintf(stderr, "\n"); } } while (0)

static bool same_bits(float a, float b) { return std::memcmp(&a, &b, 4) == 0; }
static bool same_sv(const std::vector<SV>& a, const std::vector<SV>& b) {
  if (a.size() != b.size()) return false;
  for (size_t r = 0; r < a.size(); ++r) {
    if (a[r].size() != b[r].size()) return false;
    for (size_t k = 0; k < a[r].size(); ++k)
      if (a[r][k].first != b[r][k].first || !same_bits(a[r][k].second, b[r][k].second)) return false;
  }
  return true;
}
static VVF dense(const std::vector<SV>& s, uint cols) {
  VVF d(s.size(), VF(cols, 0.0f));
  for (size_t r = 0; r < s.size(); ++r) for (const auto& e : s[r]) d[r][e.first] = e.second;
  return d;
}
static std::vector<float> flat(const VVF& v) {
  std::vector<float> f;
  for (const auto& r : v) f.insert(f.end(), r.begin(), r.end());
  return f;
}

int main(int argc, char** argv) {
  if (argc < 2) { std::fprintf(stderr, "usage: adapter_driver FASTA\n"); return 2; }
  std::vector<Fasta> fa;
  {
    std::ifstream in(argv[1]);
    std::string line, name, seq;
    while (std::getline(in, line)) {
      if (!line.empty() && line[0] == '>') { if (!seq.empty()) fa.emplace_back(name, seq); name = line.substr(1); seq.clear(); }
      else seq += line;
    }
    if (!seq.empty()) fa.emplace_back(name, seq);
  }
  const uint N = fa.size();
  CHECK(N >= 4, "need at least four sequences, got %u", N);
  try {
    // (1) alignment model: all pairs in one device call, and one pair at a time
    ProbCons cpu_a(0.01f);
    CudaProbCons gpu_a(0.01f);
    std::vector<std::vector<MP> > mp_c, mp_g;
    static_cast<Align::Model&>(cpu_a).calculate(fa, mp_c);   // the batch virtual of the base class (src/align.cpp:35-52)
    gpu_a.calculate(fa, mp_g);
    CHECK(mp_c.size() == mp_g.size(), "mp size");
    uint64_t entries = 0;
    for (uint i = 0; i < N; ++i)
      for (uint j = i; j < N; ++j) {
        CHECK(same_sv(mp_c[i][j], mp_g[i][j]), "CudaProbCons::calculate(fa): mp[%u][%u] differs", i, j);
        for (const auto& r : mp_c[i][j]) entries += r.size();
      }
    MP one_c, one_g;
    cpu_a.calculate(fa[1].seq(), fa[3].seq(), one_c);
    gpu_a.calculate(fa[1].seq(), fa[3].seq(), one_g);
    CHECK(same_sv(one_c, one_g), "CudaProbCons::calculate(seq1, seq2) differs");
    CHECK(gpu_a.threshold() == cpu_a.threshold(), "threshold()");

    // (2) fold model (no CPU class links here: RNAfold is ViennaRNA): shape of the result and batch == single
    CudaBoltzmann gpu_s(0.01f);
    std::vector<BP> bp;
    gpu_s.calculate(fa, bp);
    CHECK(bp.size() == N, "bp size");
    for (uint s = 0; s < N; ++s) {
      CHECK(bp[s].size() == fa[s].size(), "bp rows");
      for (uint i = 0; i < bp[s].size(); ++i)
        for (size_t k = 0; k < bp[s][i].size(); ++k) {
          CHECK(bp[s][i][k].first > i && bp[s][i][k].second > 0.01f && bp[s][i][k].second <= 1.0f, "bp entry out of range");
          if (k) CHECK(bp[s][i][k - 1].first < bp[s][i][k].first, "bp columns not ascending");
        }
    }
    BP bp0;
    gpu_s.calculate(fa[0].seq(), bp0);
    CHECK(same_sv(bp0, bp[0]), "CudaBoltzmann single != batch");
    // the constrained overload (src/fold.cpp:118-166): '?' everywhere = unconstrained, '.' everywhere = no pair
    BP bpq, bpx;
    gpu_s.calculate(fa[0].seq(), std::string(fa[0].size(), '?'), bpq);
    CHECK(same_sv(bpq, bp[0]), "CudaBoltzmann constrained with '?' only != unconstrained");
    gpu_s.calculate(fa[0].seq(), std::string(fa[0].size(), '.'), bpx);
    for (const auto& r : bpx) CHECK(r.empty(), "CudaBoltzmann constrained with '.' only kept a pair");

    // (3) decoders on the matrices DAFS would hand them
    const uint L1 = fa[0].size(), L2 = fa[1].size();
    const VVF p_x = dense(bp[0], L1), p_y = dense(bp[1], L2), p_z = dense(mp_c[0][1], L2);
    VVF q_x(L1, VF(L1, 0.0f)), q_z(L1, VF(L2, 0.0f));
    for (uint i = 0; i < L1; ++i) for (uint j = i + 1; j < L1; ++j) if ((i * 31 + j * 17) % 11 == 0) q_x[i][j] = 0.05f * ((i + j) % 7) - 0.1f;
    for (uint i = 0; i < L1; ++i) for (uint k = 0; k < L2; ++k) if ((i * 13 + k * 7) % 9 == 0) q_z[i][k] = 0.03f * ((i + k) % 5);
    SparseNussinov cpu_n(0.2f);
    CudaNussinov gpu_n(0.2f);
    VU ss_c, ss_g;
    const float sn_c = cpu_n.decode(2.0f, p_x, q_x, ss_c), sn_g = gpu_n.decode(2.0f, p_x, q_x, ss_g);
    CHECK(same_bits(sn_c, sn_g) && ss_c == ss_g, "CudaNussinov::decode(w,p,q): %g vs %g", sn_c, sn_g);
    std::string str_c, str_g;
    const float sp_c = cpu_n.decode(p_x, ss_c, str_c), sp_g = gpu_n.decode(p_x, ss_g, str_g);
    CHECK(same_bits(sp_c, sp_g) && ss_c == ss_g && str_c == str_g, "CudaNussinov::decode(p): %s vs %s", str_c.c_str(), str_g.c_str());
    SparseNeedlemanWunsch cpu_w(0.01f);
    CudaNeedlemanWunsch gpu_w(0.01f);
    cpu_w.initialize(p_z); gpu_w.initialize(p_z);
    VU al_c, al_g;
    const float sw_c = cpu_w.decode(p_z, q_z, al_c), sw_g = gpu_w.decode(p_z, q_z, al_g);
    CHECK(same_bits(sw_c, sw_g) && al_c == al_g, "CudaNeedlemanWunsch::decode(p,q): %g vs %g", sw_c, sw_g);
    const float sv_c = cpu_w.decode(p_z, al_c), sv_g = gpu_w.decode(p_z, al_g);
    CHECK(same_bits(sv_c, sv_g) && al_c == al_g, "CudaNeedlemanWunsch::decode(p): %g vs %g", sv_c, sv_g);

    // (4) the dual-decomposition solve of two leaves
    dafs_cuda::DDOptions o{4.0f, VF(1, 0.2f), 0.01f, 0.5f, 600};
    VU x, y, z;
    ALN a1(1), a2(1);
    const dafs_cuda::DDOutcome r = dafs_cuda::solve_by_dd(o, x, y, z, p_x, p_y, p_z, a1, a2);
    std::vector<unsigned> rx(L1), ry(L2), rz(L1);
    int it = 0; unsigned vi = 0;
    const std::vector<float> fx = flat(p_x), fy = flat(p_y), fz = flat(p_z);
    const float rs = ref_dd_solve(L1, L2, 1, 1, fx.data(), fy.data(), fz.data(), 4.0f, 0.2f, 0.01f, 0.5f, 600,
                                  rx.data(), ry.data(), rz.data(), &it, &vi);
    CHECK(same_bits(rs, r.score) && (unsigned)it == r.iters && vi == r.violated, "solve_by_dd: score %g/%g iters %d/%u violated %u/%u", rs, r.score, it, r.iters, vi, r.violated);
    CHECK(VU(rx.begin(), rx.end()) == x && VU(ry.begin(), ry.end()) == y && VU(rz.begin(), rz.end()) == z, "solve_by_dd: x / y / z differ");
    // (5) the whole pipeline, stage by stage through host containers vs in one call with device batches
    {
      std::vector<std::vector<MP> > mp2 = mp_g;
      std::vector<BP> bp2 = bp;
      VVF sim_a, sim_b;
      dafs_cuda::consistency_transform(fa, mp2, bp2, 0.25f, 0.25f, sim_a);
      std::vector<dafs_cuda::TreeNode> tree_a, tree_b;
      dafs_cuda::build_tree(sim_a, tree_a);
      VU ss_a, ss_b; ALN aln_a, aln_b;
      const float sa = dafs_cuda::align_tree(o, fa, mp2, bp2, tree_a, ss_a, aln_a);
      const float sb = dafs_cuda::align_from_sequences(o, fa, 0.01f, 0.25f, 0.25f, sim_b, tree_b, ss_b, aln_b);
      CHECK(same_bits(sa, sb) && ss_a == ss_b && aln_a == aln_b && tree_a == tree_b, "align_from_sequences differs from the stage-by-stage adapters: %g vs %g", sa, sb);
    }
    if (g_fail == 0)
      std::printf("adapters ok: %u sequences, %llu posteriors identical (ProbCons), Nussinov %.6f, NW %.6f, DD %u iterations score %.6f\n",
                  N, (unsigned long long)entries, sn_g, sw_g, r.iters, r.score);
  } catch (const std::string& e) {
    std::fprintf(stderr, "exception: %s\n", e.c_str()); return 3;
  } catch (const char* e) {
    std::fprintf(stderr, "exception: %s\n", e); return 3;
  }
  return g_fail ? 1 : 0;
}
