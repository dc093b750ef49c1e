// C++ adapters that put libdafs_cuda behind the reference's own interfaces, so
// the `dafs` driver (src/dafs.cpp) keeps every call site:
//   CudaProbCons          : Align::Model    (src/align.h:37-55)   replaces ProbCons   (src/align.cpp:54-79)
//   CudaBoltzmann         : Fold::Model     (src/fold.h:33-45)    replaces RNAfold(bl=true) (src/fold.cpp:70-116)
//   CudaNussinov          : Fold::Decoder   (src/fold.h:47-60)    replaces SparseNussinov (src/nussinov.cpp:207-413)
//   CudaNeedlemanWunsch   : Align::Decoder  (src/align.h:57-65)   replaces SparseNeedlemanWunsch (src/needleman_wunsch.cpp:198-422)
//   dafs_cuda::solve_by_dd, solve_by_dd_batch                     replace DAFS::solve_by_dd (src/dafs.cpp:1006-1295)
//   dafs_cuda::consistency_transform                              replaces src/dafs.cpp:1797-1827 of DAFS::run()
//   dafs_cuda::build_tree, align_tree                             replace DAFS::build_tree / DAFS::align (src/dafs.cpp:447-495, 1518-1537)
// This header is compiled inside the reference tree (it includes align.h and
// fold.h from there); see INTEGRATION.md for the three-line patch of
// DAFS::parse_options / DAFS::solve_by_dd.  Errors from the C ABI are re-thrown
// as std::string, which main() already catches (src/dafs.cpp:1898-1909).
#ifndef DAFS_CUDA_MODELS_H_
#define DAFS_CUDA_MODELS_H_

#include <string>
#include <vector>

#include "align.h"
#include "fold.h"
#include "typedefs.h"

class CudaProbCons : public Align::Model {
 public:
  explicit CudaProbCons(float th, int ngpu = 0);
  // one pair = a batch of one (same result as ProbCons::calculate)
  void calculate(const std::string& seq1, const std::string& seq2, MP& mp) override;
  // all i<j pairs in one device call; fills mp[i][i] with the identity exactly
  // like Align::Model::calculate (src/align.cpp:35-52)
  void calculate(const std::vector<Fasta>& fa, std::vector<std::vector<MP> >& mp) override;

 private:
  int ngpu_;
};

class CudaBoltzmann : public Fold::Model {
 public:
  explicit CudaBoltzmann(float th, int ngpu = 0);
  void calculate(const std::string& seq, BP& bp) override;
  // constrained folding (--bp-update, --bp-update1, --ipknot): src/fold.cpp:118-166
  void calculate(const std::string& seq, const std::string& str, BP& bp) override;
  void calculate(const std::vector<Fasta>& fa, std::vector<BP>& bp) override;

 private:
  int ngpu_;
};

class CudaNussinov : public Fold::Decoder {
 public:
  explicit CudaNussinov(float th) : Fold::Decoder(), th_(th) {}
  float decode(float w, const VVF& p, const VVF& q, VU& ss) override;
  float decode(const VVF& p, VU& ss, std::string& str) override;
  void make_brackets(const VU& ss, std::string& str) const override;

 private:
  float th_;
};

class CudaNeedlemanWunsch : public Align::Decoder {
 public:
  explicit CudaNeedlemanWunsch(float th) : Align::Decoder(), th_(th) {}
  void initialize(const VVF& p) override {}   // the envelope is rebuilt inside each device call
  float decode(const VVF& p, const VVF& q, VU& al) const override;
  float decode(const VVF& p, VU& al) const override;

 private:
  float th_;
};

namespace dafs_cuda {

struct DDOptions {        // the DAFS members solve_by_dd reads (src/dafs.cpp:126-135)
  float w;                // w_      (-w)
  VF th_s;                // th_s_   (-t)
  float th_a;             // th_a_   (-u)
  float eta0;             // eta0_   (--eta)
  unsigned t_max;         // t_max_  (-m)
};

struct DDOutcome { float score; unsigned iters; unsigned violated; };

// Drop-in body for DAFS::solve_by_dd: same arguments, same return value; iters and
// violated are what it logs at src/dafs.cpp:1292.
DDOutcome solve_by_dd(const DDOptions& o, VU& x, VU& y, VU& z, const VVF& p_x, const VVF& p_y,
                      const VVF& p_z, const ALN& aln1, const ALN& aln2);

// Independent problems (sibling guide-tree nodes) in one device launch.
struct DDInput { const VVF* p_x; const VVF* p_y; const VVF* p_z; unsigned n1, n2; };
struct DDOutput { VU x, y, z; DDOutcome outcome; };
void solve_by_dd_batch(const DDOptions& o, const std::vector<DDInput>& in, std::vector<DDOutput>& out);

// Drop-in for the block of DAFS::run() between the posterior calls and build_tree()
// (src/dafs.cpp:1797-1827, without the optional four-way step): transpose_mp for all
// pairs, sim_, relax_basepairing_probability (w_pct_s != 0), relax_matching_probability
// (w_pct_a != 0).  mp is what a_model_->calculate(fa_, mp_) left (mp[i][j] for i <= j);
// on return mp[i][j] is filled for all i, j, bp and sim as the reference leaves them.
void consistency_transform(const std::vector<Fasta>& fa, std::vector<std::vector<MP> >& mp, std::vector<BP>& bp,
                           float w_pct_a, float w_pct_s, VVF& sim);

// Drop-in for build_tree() + align(ss, aln, root) of DAFS::run() (src/dafs.cpp:1829-1838,
// :447-495, :1518-1537) when use_alifold_ and use_bp_update_ are off (--no-alifold): all
// guide-tree nodes whose children are finished are solved in one device wave.  tree is
// DAFS::tree_ (vector<pair<float, pair<uint, uint>>>), filled like build_tree() fills it;
// ss / aln / the return value are what align(ss, aln, tree_.size() - 1) yields.
typedef std::pair<float, std::pair<uint, uint> > TreeNode;
void build_tree(const VVF& sim, std::vector<TreeNode>& tree);
float align_tree(const DDOptions& o, const std::vector<Fasta>& fa, const std::vector<std::vector<MP> >& mp,
                 const std::vector<BP>& bp, const std::vector<TreeNode>& tree, VU& ss, ALN& aln);

// One call for DAFS::run() :1787-1838 when alifold and --bp-update are off and -f is 0: fold model, align
// model, consistency transformation, guide tree and progressive alignment with the posteriors kept in
// HBM between the stages (device batches).  th_bp = CUTOFF (:65), w_pct_a / w_pct_s = -p / -q; sim and
// tree are filled as the reference fills sim_ and tree_; ss / aln / the return value are align()'s.
float align_from_sequences(const DDOptions& o, const std::vector<Fasta>& fa, float th_bp, float w_pct_a, float w_pct_s,
                           VVF& sim, std::vector<TreeNode>& tree, VU& ss, ALN& aln);

}  // namespace dafs_cuda

#endif  // DAFS_CUDA_MODELS_H_
