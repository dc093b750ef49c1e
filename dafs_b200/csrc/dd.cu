// Host side of (3): per-problem setup (consensus-base-pair enumeration, sparse
// multiplier index sets, NW envelope), scratch layout, launch, and the C-ABI
// entry points declared in include/dafs_cuda.h.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <mutex>
#include <thread>
#include <cstdlib>
#include <cstdio>
#include <vector>

#include "common.cuh"
#include "dd_kernel.cuh"
#include "dd_pack.cuh"

namespace dafs {

static const double kCutoff = 0.01;       // CUTOFF, src/dafs.cpp:65 (a double literal)
static const int kMaxL = 16384;

// Host image of one problem: the position-independent constant block the kernel
// stages into shared memory (DDHdr + sections), and where its results go.
struct NusHost {
  DDNusHdr h{};                 // offsets into the block
  int max_col = 0;              // most candidates in one column
  size_t out = 0;               // offset in the result buffer
  std::vector<int> row_ptr_full;   // [L+1] over all entries (host only: CBP -> entry id)
  std::vector<int> row_ent_full;
  std::vector<uint32_t> ij;
};
struct ProbHost {
  NusHost sx, sy;
  int L1 = 0, L2 = 0, nb = 0, bw = 0, ncz = 0, ncbp = 0, t_max = 0, stage = 0;
  size_t z_out = 0, result = 0, trace = 0;                     // result buffer
  std::vector<char> cst;                                       // the constant block
  size_t res_bytes = 0;
  size_t cst_base = 0, res_base = 0;                           // position in the plan's buffers
  size_t cst_bytes_dev = 0;                                    // block size when the block is built on the device (cst stays empty)
};

static size_t up16(size_t x) { return (x + 15) & ~size_t(15); }

template <class T>
static size_t put(std::vector<char>& buf, const T* data, size_t n) {
  const size_t off = up16(buf.size());
  buf.resize(off + sizeof(T) * std::max<size_t>(n, 1));
  if (n) std::memcpy(buf.data() + off, data, sizeof(T) * n);
  return off;
}

struct Ent { int i, j; float p; uint8_t flag; bool cand; };

// Packs one Nussinov structure into the block: candidates (span > 2) first, in
// column order, then the span <= 2 projections; live rows / columns and the
// index maps of the compressed grid (dd_kernel.cuh).
static void pack_nus(NusHost& H, std::vector<char>& cst, int L, float w, std::vector<Ent>& ents) {
  std::stable_sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) {
    if (a.cand != b.cand) return a.cand;                       // candidates first
    return a.j != b.j ? a.j < b.j : a.i > b.i;                 // ascending j, descending i
  });
  const int n = (int)ents.size();
  int nd = 0;
  while (nd < n && ents[nd].cand) ++nd;
  std::vector<uint16_t> rowI, colJ;
  {
    std::vector<char> ri(L, 0), cj(L, 0);
    for (int e = 0; e < nd; ++e) { ri[ents[e].i] = 1; cj[ents[e].j] = 1; }
    for (int k = 0; k < L; ++k) { if (ri[k]) rowI.push_back((uint16_t)k); if (cj[k]) colJ.push_back((uint16_t)k); }
  }
  const int nI = (int)rowI.size(), nJ = (int)colJ.size();
  std::vector<uint16_t> rowmap(L + 1), colmap(L + 1), bstart(std::max(nI, 1));
  {
    int a = 0;
    for (int i = 0; i <= L; ++i) { while (a < nI && rowI[a] < i) ++a; rowmap[i] = (uint16_t)a; }
    int b = 0;                                                // colmap[j] = #live columns <= j
    for (int j = 0; j <= L; ++j) { while (b < nJ && colJ[b] <= j) ++b; colmap[j] = (uint16_t)b; }
    for (int x = 0; x < nI; ++x) bstart[x] = colmap[rowI[x]];   // first live column index with J > I[a]
  }
  std::vector<int> col_ptr(nJ + 1, 0), row_ptr(nI + 1, 0), row_ent(std::max(nd, 1)), ent_c(std::max(nd, 1));
  std::vector<uint32_t> ab(std::max(nd, 1)), ij(std::max(n, 1));
  std::vector<float> pv(std::max(n, 1));
  std::vector<uint8_t> fl(std::max(n, 1));
  for (int e = 0; e < n; ++e) {
    ij[e] = (uint32_t)ents[e].i | ((uint32_t)ents[e].j << 16);
    pv[e] = ents[e].p; fl[e] = ents[e].flag;
  }
  int max_col = 0;
  for (int e = 0; e < nd; ++e) {
    const int a = rowmap[ents[e].i], b = (int)colmap[ents[e].j] - 1;
    ab[e] = (uint32_t)a | ((uint32_t)b << 16);
    // live row a in the high half; low half: 1 + (last live column <= i-1), 0 if none
    ent_c[e] = (a << 16) | (ents[e].i >= 1 ? (int)colmap[ents[e].i - 1] : 0);
    col_ptr[b + 1]++; row_ptr[a + 1]++;
  }
  for (int b = 0; b < nJ; ++b) { max_col = std::max(max_col, col_ptr[b + 1]); col_ptr[b + 1] += col_ptr[b]; }
  for (int a = 0; a < nI; ++a) row_ptr[a + 1] += row_ptr[a];
  {
    std::vector<int> fill(row_ptr.begin(), row_ptr.end() - 1);
    for (int e = 0; e < nd; ++e) row_ent[fill[ab[e] & 0xffffu]++] = e;   // ascending b inside a row
  }
  // host-only full row lists (all entries) for CBP -> entry id
  H.row_ptr_full.assign(L + 1, 0);
  for (int e = 0; e < n; ++e) H.row_ptr_full[ents[e].i + 1]++;
  for (int k = 0; k < L; ++k) H.row_ptr_full[k + 1] += H.row_ptr_full[k];
  H.row_ent_full.assign(std::max(n, 1), 0);
  {
    std::vector<int> fill(H.row_ptr_full.begin(), H.row_ptr_full.end() - 1);
    for (int e = 0; e < n; ++e) H.row_ent_full[fill[ents[e].i]++] = e;
  }
  H.ij = ij;
  DDNusHdr& h = H.h;
  h.L = L; h.n = n; h.nd = nd; h.nI = nI; h.nJ = nJ; h.w = w;
  H.max_col = max_col;
  h.col_ptr = (uint32_t)put(cst, col_ptr.data(), col_ptr.size());
  h.row_ptr = (uint32_t)put(cst, row_ptr.data(), row_ptr.size());
  h.row_ent = (uint32_t)put(cst, row_ent.data(), (size_t)nd);
  h.ent_ab = (uint32_t)put(cst, ab.data(), (size_t)nd);
  h.ent_c = (uint32_t)put(cst, ent_c.data(), (size_t)nd);
  h.ent_ij = (uint32_t)put(cst, ij.data(), (size_t)n);
  h.ent_p = (uint32_t)put(cst, pv.data(), (size_t)n);
  h.ent_flag = (uint32_t)put(cst, fl.data(), (size_t)n);
  h.rowI = (uint32_t)put(cst, rowI.data(), rowI.size());
  h.colJ = (uint32_t)put(cst, colJ.data(), colJ.size());
  h.bstart = (uint32_t)put(cst, bstart.data(), (size_t)nI);
  h.rowmap = (uint32_t)put(cst, rowmap.data(), rowmap.size());
  h.colmap = (uint32_t)put(cst, colmap.data(), colmap.size());
}

// Entries of one structure: decoder candidates (w*(p-th) > 0 with q = 0, span
// > 2: src/nussinov.cpp:234-237) plus the projection c_x of the CBPs
// (src/dafs.cpp:1037).  Nothing else can ever get a non-zero multiplier:
// q changes only for c_x entries and for decoded pairs (dafs.cpp:1121-1150),
// and a decoded pair had s > 0.  A c_x entry with span > 2 is a candidate too
// (its multiplier can turn s positive).
static void build_nus(NusHost& H, std::vector<char>& cst, size_t& res, int L, const float* p, float w,
                      float th, const std::vector<std::pair<int, int>>& proj /* sorted unique (i,j) */) {
  std::vector<Ent> ents;
  size_t pc = 0;
  for (int i = 0; i < L; ++i)
    for (int j = i + 1; j < L; ++j) {
      const float pij = p[(size_t)i * L + j];
      bool inproj = false;
      while (pc < proj.size() && (proj[pc].first < i || (proj[pc].first == i && proj[pc].second < j))) ++pc;
      if (pc < proj.size() && proj[pc].first == i && proj[pc].second == j) inproj = true;
      const bool span = (j - i > 2);
      const float s0 = w * (pij - th);
      if ((span && s0 > 0.0f) || inproj) ents.push_back({i, j, pij, (uint8_t)(inproj ? 1 : 0), span});
    }
  pack_nus(H, cst, L, w, ents);
  H.out = up16(res); res = H.out + sizeof(uint32_t) * L;
}

// entry id of (i, j) in a structure built by pack_nus (must exist)
static int find_ent(const NusHost& H, int i, int j) {
  int lo = H.row_ptr_full[i], hi = H.row_ptr_full[i + 1];
  for (int k = lo; k < hi; ++k)
    if ((int)(H.ij[H.row_ent_full[k]] >> 16) == j) return H.row_ent_full[k];
  return -1;
}

// SparseNeedlemanWunsch::initialize, src/needleman_wunsch.cpp:198-253
static void build_envelope(int L1, int L2, float th, const float* p, std::vector<uint2>& env) {
  env.assign(L1 + 1, make_uint2(0, 0));
  for (int i = 1; i <= L1; ++i) {
    const float* row = p + (size_t)(i - 1) * L2;
    for (int k = 1; k != L2 + 1; ++k)
      if (row[k - 1] - th >= 0.0) {
        env[i - 1].x = std::min(env[i - 1].x, (unsigned)(k - 1));
        env[i].x = (unsigned)k;
        break;
      }
    if (env[i].x == 0) { env[i] = env[i - 1]; continue; }
    for (int k = L2; k != 0; --k)
      if (row[k - 1] - th >= 0.0) {
        env[i - 1].y = std::max(env[i - 1].y, (unsigned)(k - 1));
        env[i].y = (unsigned)k;
        break;
      }
  }
  env[L1].y = (unsigned)L2;
  { unsigned v = (unsigned)L2; for (int i = L1; i != 0; --i) env[i].x = v = std::min(v, env[i].x); }
  { unsigned v = 0; for (int i = 0; i != L1 + 1; ++i) env[i].y = v = std::max(v, env[i].y); }
  for (int i = 1; i != L1 + 1; ++i)
    if (env[i - 1].y < env[i].x) env[i].x = env[i - 1].y;
}

struct Cbp { int i, j, k, l; };

// src/dafs.cpp:1016-1061, walking only the entries that pass the CUTOFF gates
// (same lexicographic (i, j, k, l) order as the reference's four nested loops).
static void enumerate_cbp(const dafs_dd_problem& q, std::vector<Cbp>& cbp) {
  const int L1 = q.L1, L2 = q.L2;
  std::vector<int> xptr(L1 + 1, 0), yptr(L2 + 1, 0), zptr(L1 + 1, 0);
  std::vector<int> xcol, ycol, zcol;
  for (int i = 0; i < L1; ++i) {
    for (int j = i + 1; j < L1; ++j) if (q.p_x[(size_t)i * L1 + j] > kCutoff) xcol.push_back(j);
    xptr[i + 1] = (int)xcol.size();
    for (int k = 0; k < L2; ++k) if (q.p_z[(size_t)i * L2 + k] > kCutoff) zcol.push_back(k);
    zptr[i + 1] = (int)zcol.size();
  }
  for (int k = 0; k < L2; ++k) {
    for (int l = k + 1; l < L2; ++l) if (q.p_y[(size_t)k * L2 + l] > kCutoff) ycol.push_back(l);
    yptr[k + 1] = (int)ycol.size();
  }
  const unsigned N1 = (unsigned)q.N1, N2 = (unsigned)q.N2;
  for (int i = 0; i + 1 < L1; ++i)
    for (int a = xptr[i]; a < xptr[i + 1]; ++a) {
      const int j = xcol[a];
      const float pxv = q.p_x[(size_t)i * L1 + j];
      for (int b = zptr[i]; b < zptr[i + 1]; ++b) {
        const int k = zcol[b];
        if (k + 1 >= L2) continue;
        const float pz1 = q.p_z[(size_t)i * L2 + k];
        for (int c = yptr[k]; c < yptr[k + 1]; ++c) {
          const int l = ycol[c];
          const float pz2 = q.p_z[(size_t)j * L2 + l];
          if (!(pz2 > kCutoff)) continue;
          const float p = (N1 * pxv + N2 * q.p_y[(size_t)k * L2 + l]) / (N1 + N2);     // :1030
          const float qq = (pz1 + pz2) / 2;                                            // :1031
          if (p - q.min_th_s > 0.0 && q.w * (p - q.min_th_s) + (qq - q.th_a) > 0.0)    // :1032
            cbp.push_back({i, j, k, l});
        }
      }
    }
}

static int build_problem(const dafs_dd_problem& q, bool want_trace, ProbHost& H) {
  const int L1 = q.L1, L2 = q.L2;
  H.L1 = L1; H.L2 = L2; H.t_max = q.t_max;
  std::vector<Cbp> cbp;
  enumerate_cbp(q, cbp);
  H.ncbp = (int)cbp.size();
  std::vector<std::pair<int, int>> cx, cy, cz;
  cx.reserve(cbp.size()); cy.reserve(cbp.size()); cz.reserve(2 * cbp.size());
  for (const Cbp& c : cbp) {
    cx.emplace_back(c.i, c.j); cy.emplace_back(c.k, c.l);
    cz.emplace_back(c.i, c.k); cz.emplace_back(c.j, c.l);
  }
  auto uniq = [](std::vector<std::pair<int, int>>& v) {
    std::sort(v.begin(), v.end());
    v.erase(std::unique(v.begin(), v.end()), v.end());
  };
  uniq(cx); uniq(cy); uniq(cz);
  const unsigned N1 = (unsigned)q.N1, N2 = (unsigned)q.N2;
  const float wx = q.w * 2 * N1 / (N1 + N2), wy = q.w * 2 * N2 / (N1 + N2);   // dafs.cpp:1091-1092
  size_t res = 0;
  H.cst.assign(sizeof(DDHdr), 0);
  build_nus(H.sx, H.cst, res, L1, q.p_x, wx, q.th_s, cx);
  build_nus(H.sy, H.cst, res, L2, q.p_y, wy, q.th_s, cy);
  // z: the aligner only ever evaluates cells inside the envelope
  // (needleman_wunsch.cpp:262-264), so p_z is kept as a dense band over it; the
  // multipliers q_z can leave 0 only on c_z, kept as a short list per row with the
  // band position of each cell.
  std::vector<uint2> env;
  build_envelope(L1, L2, q.th_a, q.p_z, env);
  std::vector<int> boff(L1 + 2, 0);
  std::vector<float> pb;
  for (int r = 1; r <= L1; ++r) {
    boff[r] = (int)pb.size();
    const int ef = std::max(1, (int)env[r].x), el = (int)env[r].y;
    for (int k = ef; k <= el; ++k) pb.push_back(q.p_z[(size_t)(r - 1) * L2 + (k - 1)]);
    H.bw = std::max(H.bw, el - ef + 1);
  }
  boff[L1 + 1] = (int)pb.size();
  H.nb = (int)pb.size();
  std::vector<int> cz_row_ptr(L1 + 1, 0), cz_bidx(cz.size());
  std::vector<uint16_t> cz_col(cz.size());
  for (size_t c = 0; c < cz.size(); ++c) {
    const int i = cz[c].first, k = cz[c].second;               // 0-based; row r = i+1, column k+1
    ++cz_row_ptr[i + 1];
    cz_col[c] = (uint16_t)k;
    const int ef = std::max(1, (int)env[i + 1].x), el = (int)env[i + 1].y;
    cz_bidx[c] = (k + 1 >= ef && k + 1 <= el) ? boff[i + 1] + (k + 1 - ef) : -1;
  }
  for (int i = 0; i < L1; ++i) cz_row_ptr[i + 1] += cz_row_ptr[i];
  H.ncz = (int)cz.size();
  DDHdr h{};
  h.z_env = (uint32_t)put(H.cst, env.data(), env.size());
  h.z_boff = (uint32_t)put(H.cst, boff.data(), boff.size());
  h.z_pb = (uint32_t)put(H.cst, pb.data(), pb.size());
  h.cz_row_ptr = (uint32_t)put(H.cst, cz_row_ptr.data(), cz_row_ptr.size());
  h.cz_col = (uint32_t)put(H.cst, cz_col.data(), cz_col.size());
  h.cz_bidx = (uint32_t)put(H.cst, cz_bidx.data(), cz_bidx.size());
  std::vector<int4> cb(cbp.size());
  auto zid = [&](int i, int k) {
    return (int)(std::lower_bound(cz.begin(), cz.end(), std::make_pair(i, k)) - cz.begin());
  };
  for (size_t u = 0; u < cbp.size(); ++u)
    cb[u] = make_int4(find_ent(H.sx, cbp[u].i, cbp[u].j), find_ent(H.sy, cbp[u].k, cbp[u].l),
                      zid(cbp[u].i, cbp[u].k), zid(cbp[u].j, cbp[u].l));
  H.cst.resize(up16(H.cst.size()));
  h.stage = (int)H.cst.size();                 // cbp stays in HBM (dd_kernel.cuh)
  H.stage = h.stage;
  h.cbp = (uint32_t)put(H.cst, cb.data(), cb.size());
  H.cst.resize(up16(H.cst.size()));
  h.L1 = L1; h.L2 = L2; h.nb = H.nb; h.ncz = H.ncz; h.ncbp = H.ncbp; h.t_max = q.t_max;
  h.bytes = (int)H.cst.size();
  h.th_s = q.th_s; h.th_a = q.th_a; h.eta0 = q.eta0;
  h.x = H.sx.h; h.y = H.sy.h;
  H.sx.row_ptr_full = std::vector<int>(); H.sx.row_ent_full = std::vector<int>(); H.sx.ij = std::vector<uint32_t>();
  H.sy.row_ptr_full = std::vector<int>(); H.sy.row_ent_full = std::vector<int>(); H.sy.ij = std::vector<uint32_t>();
  std::memcpy(H.cst.data(), &h, sizeof h);
  H.z_out = up16(res); res = H.z_out + sizeof(uint32_t) * L1;
  H.result = up16(res); res = H.result + 64;
  H.trace = up16(res);
  if (want_trace) res = H.trace + sizeof(float) * 3 * (size_t)std::max(q.t_max, 0);
  H.res_bytes = up16(res);
  return DAFS_CUDA_OK;
}

// Offsets of every section of a problem's constant block from its counts alone: the same sequence of
// put() calls as pack_nus() / build_problem() above (16-byte alignment, at least one element each).
static void layout_nus(size_t& cur, DDNusHdr& h, int L, int n, int nd, int nI, int nJ, float w) {
  auto sect = [&](size_t elem, size_t cnt) { const size_t off = up16(cur); cur = off + elem * std::max<size_t>(cnt, 1); return (uint32_t)off; };
  h = DDNusHdr{};
  h.L = L; h.n = n; h.nd = nd; h.nI = nI; h.nJ = nJ; h.w = w;
  h.col_ptr = sect(4, (size_t)nJ + 1); h.row_ptr = sect(4, (size_t)nI + 1); h.row_ent = sect(4, nd);
  h.ent_ab = sect(4, nd); h.ent_c = sect(4, nd); h.ent_ij = sect(4, n); h.ent_p = sect(4, n); h.ent_flag = sect(1, n);
  h.rowI = sect(2, nI); h.colJ = sect(2, nJ); h.bstart = sect(2, nI); h.rowmap = sect(2, (size_t)L + 1); h.colmap = sect(2, (size_t)L + 1);
}

static void layout_problem(const dafs_dd_problem& q, const PackCounts& c, bool want_trace, ProbHost& H, DDHdr& h) {
  const int L1 = q.L1, L2 = q.L2;
  const unsigned N1 = (unsigned)q.N1, N2 = (unsigned)q.N2;
  const float wx = q.w * 2 * N1 / (N1 + N2), wy = q.w * 2 * N2 / (N1 + N2);   // dafs.cpp:1091-1092
  H = ProbHost();
  H.L1 = L1; H.L2 = L2; H.t_max = q.t_max; H.ncbp = c.ncbp; H.nb = c.nb; H.bw = c.bw; H.ncz = c.ncz;
  h = DDHdr{};
  size_t cur = sizeof(DDHdr), res = 0;
  layout_nus(cur, H.sx.h, L1, c.n_x, c.nd_x, c.nI_x, c.nJ_x, wx);
  H.sx.max_col = c.maxcol_x; H.sx.out = up16(res); res = H.sx.out + sizeof(uint32_t) * L1;
  layout_nus(cur, H.sy.h, L2, c.n_y, c.nd_y, c.nI_y, c.nJ_y, wy);
  H.sy.max_col = c.maxcol_y; H.sy.out = up16(res); res = H.sy.out + sizeof(uint32_t) * L2;
  auto sect = [&](size_t elem, size_t cnt) { const size_t off = up16(cur); cur = off + elem * std::max<size_t>(cnt, 1); return (uint32_t)off; };
  h.z_env = sect(8, (size_t)L1 + 1); h.z_boff = sect(4, (size_t)L1 + 2); h.z_pb = sect(4, c.nb);
  h.cz_row_ptr = sect(4, (size_t)L1 + 1); h.cz_col = sect(2, c.ncz); h.cz_bidx = sect(4, c.ncz);
  cur = up16(cur);
  h.stage = (int)cur; H.stage = h.stage;
  h.cbp = sect(16, c.ncbp);
  cur = up16(cur);
  h.L1 = L1; h.L2 = L2; h.nb = c.nb; h.ncz = c.ncz; h.ncbp = c.ncbp; h.t_max = q.t_max;
  h.bytes = (int)cur;
  h.th_s = q.th_s; h.th_a = q.th_a; h.eta0 = q.eta0;
  h.x = H.sx.h; h.y = H.sy.h;
  H.cst_bytes_dev = cur;
  H.z_out = up16(res); res = H.z_out + sizeof(uint32_t) * L1;
  H.result = up16(res); res = H.result + 64;
  H.trace = up16(res);
  if (want_trace) res = H.trace + sizeof(float) * 3 * (size_t)std::max(q.t_max, 0);
  H.res_bytes = up16(res);
}

}  // namespace dafs

// ---- plan ---------------------------------------------------------------------
struct dafs_dd_plan {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int nprob = 0;
  bool want_trace = false;
  std::vector<dafs::ProbHost> host;       // offsets (cst blocks are released after upload)
  char* d_cst = nullptr; size_t cst_bytes = 0;
  char* d_res = nullptr; size_t res_bytes = 0;
  dafs::DDProbRef* d_prob = nullptr;
  int32_t* d_order = nullptr;
  int* d_queue = nullptr;
  char* d_ws = nullptr;
  dafs::DDLayout lay{};
  int grid = 0, smem = 0, ctas_per_sm = 0;
  bool all_smem = false;
  int team = 1;                            // warps per Nussinov structure (dd_kernel<SMEM, T>)
  cudaStream_t poll_stream = nullptr;      // dafs_cuda_dd_plan_poll: copies under the running kernel
  bool own_stream = true, own_poll = true; // false: lent by the caller (common.cuh: g_dd_stream)
  char* poll_stage = nullptr;              // pinned, res_bytes
  int64_t total_cbp = 0, total_entries = 0;
  bool launched = false;
  bool packed_on_device = false;
  float last_ms = 0;
};

namespace dafs {

static void free_dd_plan(dafs_dd_plan* p) {
  dev_free(p->d_cst); dev_free(p->d_res); dev_free(p->d_prob); dev_free(p->d_order);
  dev_free(p->d_queue); dev_free(p->d_ws);
  if (p->ev0) cudaEventDestroy(p->ev0);
  if (p->ev1) cudaEventDestroy(p->ev1);
  if (p->poll_stream && p->own_poll) cudaStreamDestroy(p->poll_stream);
  if (p->poll_stage) pinned_free(p->poll_stage);
  if (p->stream && p->own_stream) cudaStreamDestroy(p->stream);
}

struct DDMaxima { int L1 = 1, L2 = 1, nx = 1, ny = 1, nb = 1, bw = 1, ncz = 1, ncbp = 1, col = 0, cst = 16, nIx = 1, nJx = 1, nIy = 1, nJy = 1; };

// Scratch layout from the batch maxima: items go to shared memory in priority
// order (enum order in dd_kernel.cuh) while they fit, the rest to the per-CTA
// global workspace.
static void plan_layout(DDLayout& lay, const DDMaxima& m, size_t smem_budget, int team, int* smem_bytes, size_t* ws_bytes) {
  lay.lp1 = (m.nJx + 2) & ~1;      // even stride: lane*lp + (B0 + s - lane) hits 32 distinct banks
  lay.lp2 = (m.nJy + 2) & ~1;
  lay.trs1 = m.nJx; lay.trs2 = m.nJy;
  lay.ts = (m.bw + 31) / 32;
  { const char* e = getenv("DAFS_DD_NW_CHUNKED"); lay.nw_chunked = (e && atoi(e) > 0) ? 1 : 0; }
  lay.tr16 = (m.col > 120) ? 1u : 0u;
  const size_t trb = lay.tr16 ? 2 : 1;
  size_t sz[DD_NITEMS];
  sz[DD_CST] = (size_t)m.cst;
  sz[DD_Q_X] = 4 * (size_t)m.nx; sz[DD_Q_Y] = 4 * (size_t)m.ny; sz[DD_Q_Z] = 4 * (size_t)m.nb; sz[DD_Q_CZ] = 4 * (size_t)m.ncz;
  sz[DD_ES_X] = 4 * (size_t)m.nx; sz[DD_ES_Y] = 4 * (size_t)m.ny;
  sz[DD_ROW_X] = 4 * (size_t)32 * lay.lp1 * team; sz[DD_ROW_Y] = 4 * (size_t)32 * lay.lp2 * team;   // one wave per team warp
  sz[DD_BROW_X] = 4 * (size_t)lay.lp1 * team; sz[DD_BROW_Y] = 4 * (size_t)lay.lp2 * team;
  sz[DD_BROW_Z] = 4 * 2 * (size_t)(m.L2 + 2);
  sz[DD_TRZ] = 8 * (size_t)(m.L1 + 1) * lay.ts;
  sz[DD_TR_X] = trb * ((size_t)m.nIx * m.nJx + 1);
  sz[DD_TR_Y] = trb * ((size_t)m.nIy * m.nJy + 1);
  sz[DD_SS_X] = 4 * (size_t)m.L1; sz[DD_SS_Y] = 4 * (size_t)m.L2; sz[DD_ZRES] = 4 * (size_t)m.L1;
  sz[DD_STK_X] = 4 * (size_t)(m.L1 / 2 + 4); sz[DD_STK_Y] = 4 * (size_t)(m.L2 / 2 + 4);
  sz[DD_TAL_X] = 4 * (size_t)m.nx; sz[DD_TAL_Y] = 4 * (size_t)m.ny; sz[DD_TAL_Z] = 4 * (size_t)m.ncz;
  sz[DD_SW] = 4 * (size_t)m.ncbp;
  size_t s_off = 0, g_off = 0;
  lay.in_smem = 0;
  for (int k = 0; k < DD_NITEMS; ++k) {
    const size_t b = up16(std::max<size_t>(sz[k], 16));
    if (s_off + b <= smem_budget) { lay.off[k] = (uint32_t)s_off; s_off += b; lay.in_smem |= 1u << k; }
    else { lay.off[k] = (uint32_t)g_off; g_off += b; }
  }
  *smem_bytes = (int)s_off;
  *ws_bytes = up16(g_off) + 256;
  if (getenv("DAFS_DD_DEBUG_LAYOUT")) {
    static const char* names[DD_NITEMS] = {"CST", "Q_X", "Q_Y", "Q_Z", "Q_CZ", "ES_X", "ES_Y", "ROW_X", "ROW_Y", "BROW_X", "BROW_Y",
                                           "BROW_Z", "TRZ", "TR_X", "TR_Y", "SS_X", "SS_Y", "ZRES", "STK_X", "STK_Y",
                                           "TAL_X", "TAL_Y", "TAL_Z", "SW"};
    for (int k = 0; k < DD_NITEMS; ++k) fprintf(stderr, "dd layout %-7s %8zu B %s\n", names[k], sz[k], (lay.in_smem >> k & 1) ? "smem" : "hbm");
    fprintf(stderr, "dd layout total smem %zu B; nx %d ny %d nb %d ncz %d ncbp %d nI/nJ x %d/%d y %d/%d\n", s_off, m.nx, m.ny, m.nb, m.ncz, m.ncbp, m.nIx, m.nJx, m.nIy, m.nJy);
  }
}

}  // namespace dafs

using dafs::fail;

namespace dafs {
static const void* dd_kernel_fn(bool smem, int team) {
  switch (team) {
    case 3: return smem ? (const void*)dd_kernel<true, 3> : (const void*)dd_kernel<false, 3>;
    case 2: return smem ? (const void*)dd_kernel<true, 2> : (const void*)dd_kernel<false, 2>;
    default: return smem ? (const void*)dd_kernel<true, 1> : (const void*)dd_kernel<false, 1>;
  }
}
// The opt-in limit is raised to the device maximum once per kernel and device and then left
// alone: launches of different sizes are in flight at the same time (dafs_cuda_progressive),
// and changing the attribute of a function that is running makes the next launch wait for it.
static cudaError_t dd_allow_smem(const void* fn, bool smem, int team, int device, size_t optin) {
  static std::mutex mu;
  static bool done[64][2][4] = {};
  std::lock_guard<std::mutex> g(mu);
  bool& d = done[device & 63][smem ? 1 : 0][team & 3];
  if (d) return cudaSuccess;
  const cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)optin - 1024);   // the kernel has some static shared memory too
  if (e == cudaSuccess) d = true;
  return e;
}
}  // namespace dafs

namespace dafs {
size_t dd_pack_bytes(int nprob, const dafs_dd_problem* prob) {
  size_t sc = 0;
  for (int k = 0; k < nprob; ++k) sc += pack_scratch_bytes(prob[k].L1, prob[k].L2);
  return sc + pack_up(sizeof(PackProb) * nprob) + pack_up(sizeof(PackCounts) * nprob) + pack_up(sizeof(DDHdr) * nprob);
}
}  // namespace dafs

// device_inputs: p_x / p_y / p_z of every problem are DEVICE pointers and the constant blocks are
// built by dd_pack_count / dd_pack_fill (dd_pack.cuh) instead of build_problem() on the host cores.
static int dd_plan_create_impl(int nprob, const dafs_dd_problem* prob, int want_trace, bool device_inputs,
                               dafs_dd_plan** out) {
  using namespace dafs;
  if (!out) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  *out = nullptr;
  if (nprob < 0 || (nprob > 0 && !prob)) return fail(DAFS_CUDA_EINVAL, "bad nprob / prob");
  for (int k = 0; k < nprob; ++k) {
    const dafs_dd_problem& q = prob[k];
    if (q.L1 < 1 || q.L2 < 1 || q.L1 > kMaxL || q.L2 > kMaxL)
      return fail(DAFS_CUDA_EINVAL, "problem %d: L1=%d L2=%d outside [1, %d]", k, q.L1, q.L2, kMaxL);
    if (!q.p_x || !q.p_y || !q.p_z) return fail(DAFS_CUDA_EINVAL, "problem %d: NULL matrix", k);
    if (q.N1 < 1 || q.N2 < 1 || q.t_max < 0) return fail(DAFS_CUDA_EINVAL, "problem %d: bad N1/N2/t_max", k);
    // q_z is kept only on c_z: with a negative step size the reference's
    // max(0, q - eta*(1-t)) would also grow multipliers of every aligned cell.
    if (q.eta0 < 0.0f) return fail(DAFS_CUDA_ENOTSUP, "problem %d: negative --eta is not supported", k);
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)");
  }
  dafs_dd_plan* p = new dafs_dd_plan();
  auto bail = [&](int rc) { free_dd_plan(p); delete p; return rc; };
#define PLAN_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return bail(fail(e_ == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM : DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_))); } while (0)
  PLAN_TRY(cudaGetDevice(&p->device));
  p->nprob = nprob; p->want_trace = want_trace != 0;
  if (g_dd_stream) { p->stream = g_dd_stream; p->own_stream = false; }
  else PLAN_TRY(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
  if (g_dd_poll_stream) { p->poll_stream = g_dd_poll_stream; p->own_poll = false; }
  PLAN_TRY(cudaEventCreate(&p->ev0));
  PLAN_TRY(cudaEventCreate(&p->ev1));
  PLAN_TRY(dev_malloc(&p->d_queue, sizeof(int)));
  if (nprob == 0) { *out = p; return DAFS_CUDA_OK; }

  p->host.resize(nprob);
  p->packed_on_device = device_inputs;
  // device path: scratch + problem descriptors, count kernel, totals back (64 bytes per problem)
  char* d_scratch = nullptr; PackProb* d_pp = nullptr; PackCounts* d_counts = nullptr; DDHdr* d_hdr = nullptr;
  PackProb* c_pp_ = nullptr; PackCounts* c_counts_ = nullptr; DDHdr* c_hdr_ = nullptr;   // carved from the scratch block
  std::vector<PackProb> pp;
  std::vector<DDHdr> hdr;
  auto pack_free = [&]() { dev_free(d_scratch); dev_free(d_pp); dev_free(d_counts); dev_free(d_hdr); d_scratch = nullptr; d_pp = nullptr; d_counts = nullptr; d_hdr = nullptr; };
  auto bail_pack = [&](int rc) { pack_free(); return bail(rc); };
#define PACK_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return bail_pack(fail(e_ == cudaErrorMemoryAllocation ? DAFS_CUDA_ENOMEM : DAFS_CUDA_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_))); } while (0)
  if (device_inputs) {
    pp.resize(nprob); hdr.resize(nprob);
    size_t sc = 0;
    std::vector<size_t> sc_off(nprob);
    for (int k = 0; k < nprob; ++k) { sc_off[k] = sc; sc += pack_scratch_bytes(prob[k].L1, prob[k].L2); }
    const size_t o_pp = sc, o_counts = o_pp + pack_up(sizeof(PackProb) * nprob), o_hdr = o_counts + pack_up(sizeof(PackCounts) * nprob);
    const size_t need = o_hdr + pack_up(sizeof(DDHdr) * nprob);
    char* blockp = nullptr;
    if (g_dd_pack_lend.base && g_dd_pack_lend.cap >= need) blockp = g_dd_pack_lend.base;      // lent by the progressive stage
    else { PACK_TRY(dev_malloc(&d_scratch, need)); blockp = d_scratch; }
    char* const scratch0 = blockp;
    d_pp = nullptr; d_counts = nullptr; d_hdr = nullptr;                                         // carved, not owned
    PackProb* const c_pp = reinterpret_cast<PackProb*>(blockp + o_pp);
    PackCounts* const c_counts = reinterpret_cast<PackCounts*>(blockp + o_counts);
    DDHdr* const c_hdr = reinterpret_cast<DDHdr*>(blockp + o_hdr);
    c_pp_ = c_pp; c_counts_ = c_counts; c_hdr_ = c_hdr;
    for (int k = 0; k < nprob; ++k) {
      const dafs_dd_problem& q = prob[k];
      PackProb& P = pp[k];
      const unsigned N1 = (unsigned)q.N1, N2 = (unsigned)q.N2;
      P.L1 = q.L1; P.L2 = q.L2; P.N1 = N1; P.N2 = N2; P.p_x = q.p_x; P.p_y = q.p_y; P.p_z = q.p_z;
      P.w = q.w; P.wx = q.w * 2 * N1 / (N1 + N2); P.wy = q.w * 2 * N2 / (N1 + N2);
      P.th_s = q.th_s; P.min_th_s = q.min_th_s; P.th_a = q.th_a;
      P.scratch = scratch0 + sc_off[k]; P.block = nullptr; P.hdr = nullptr;
    }
    std::vector<PackCounts> counts(nprob);
    PACK_TRY(cudaMemcpyAsync(c_pp_, pp.data(), sizeof(PackProb) * nprob, cudaMemcpyHostToDevice, p->stream));
    int maxL = 1;
    for (int k = 0; k < nprob; ++k) maxL = std::max(maxL, std::max(prob[k].L1, prob[k].L2));
    const dim3 rgrid((maxL + DD_PACK_ROWS - 1) / DD_PACK_ROWS, nprob);
    const size_t env_smem = (size_t)(maxL + 1) * sizeof(uint2) <= 96 * 1024 ? (size_t)(maxL + 1) * sizeof(uint2) : 0;
    if (env_smem > 48 * 1024) PACK_TRY(cudaFuncSetAttribute(dd_pack_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)env_smem));
    dd_pack_gates<<<rgrid, DD_PACK_THREADS, 0, p->stream>>>(c_pp_);
    dd_pack_cbp<<<rgrid, DD_PACK_THREADS, 0, p->stream>>>(c_pp_);
    dd_pack_classify<<<rgrid, DD_PACK_THREADS, 0, p->stream>>>(c_pp_);
    dd_pack_scan<<<nprob, DD_PACK_THREADS, env_smem, p->stream>>>(c_pp_, c_counts_);
    PACK_TRY(cudaGetLastError());
    PACK_TRY(cudaMemcpyAsync(counts.data(), c_counts_, sizeof(PackCounts) * nprob, cudaMemcpyDeviceToHost, p->stream));
    PACK_TRY(cudaStreamSynchronize(p->stream));
    for (int k = 0; k < nprob; ++k) {
      if (counts[k].n_x > 65535 * 8 || counts[k].ncz > (1 << 28)) return bail_pack(fail(DAFS_CUDA_EINVAL, "problem %d is too dense", k));
      layout_problem(prob[k], counts[k], p->want_trace, p->host[k], hdr[k]);
    }
  } else {
    // per-problem setup on the host cores (independent problems)
    std::atomic<int> next(0);
    const int nt = (int)std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)nprob));
    auto work = [&]() { for (;;) { const int k = next.fetch_add(1); if (k >= nprob) break; build_problem(prob[k], p->want_trace, p->host[k]); } };
    if (nt == 1) work();
    else { std::vector<std::thread> th; for (int k = 0; k < nt; ++k) th.emplace_back(work); for (auto& t : th) t.join(); }
  }
  DDMaxima m;
  size_t cst = 0, res = 0;
  for (auto& H : p->host) {
    H.cst_base = cst; cst += device_inputs ? H.cst_bytes_dev : H.cst.size(); cst = (cst + 255) & ~size_t(255);
    H.res_base = res; res += H.res_bytes;
    m.L1 = std::max(m.L1, H.L1); m.L2 = std::max(m.L2, H.L2);
    m.nx = std::max(m.nx, H.sx.h.n); m.ny = std::max(m.ny, H.sy.h.n);
    m.nIx = std::max(m.nIx, H.sx.h.nI); m.nJx = std::max(m.nJx, H.sx.h.nJ);
    m.nIy = std::max(m.nIy, H.sy.h.nI); m.nJy = std::max(m.nJy, H.sy.h.nJ);
    m.nb = std::max(m.nb, H.nb); m.bw = std::max(m.bw, H.bw); m.ncz = std::max(m.ncz, H.ncz); m.ncbp = std::max(m.ncbp, H.ncbp);
    m.col = std::max(m.col, std::max(H.sx.max_col, H.sy.max_col));
    m.cst = std::max(m.cst, H.stage);
    p->total_cbp += H.ncbp; p->total_entries += H.sx.h.n + H.sy.h.n + H.nb;
  }
  p->cst_bytes = cst; p->res_bytes = res;
  PLAN_TRY(dev_malloc(&p->d_cst, std::max<size_t>(cst, 256)));
  PLAN_TRY(dev_malloc(&p->d_res, std::max<size_t>(res, 256)));
  PLAN_TRY(dev_malloc(&p->d_prob, sizeof(DDProbRef) * nprob));
  PLAN_TRY(dev_malloc(&p->d_order, sizeof(int32_t) * nprob));
  if (device_inputs) {
    // blocks start as zeros (the host packer's padding bytes are zeros), then every section is written in place
    PACK_TRY(cudaMemsetAsync(p->d_cst, 0, std::max<size_t>(cst, 256), p->stream));
    for (int k = 0; k < nprob; ++k) { pp[k].block = p->d_cst + p->host[k].cst_base; pp[k].hdr = c_hdr_ + k; }
    PACK_TRY(cudaMemcpyAsync(c_pp_, pp.data(), sizeof(PackProb) * nprob, cudaMemcpyHostToDevice, p->stream));
    PACK_TRY(cudaMemcpyAsync(c_hdr_, hdr.data(), sizeof(DDHdr) * nprob, cudaMemcpyHostToDevice, p->stream));
    {
      int maxL = 1;
      for (int k = 0; k < nprob; ++k) maxL = std::max(maxL, std::max(prob[k].L1, prob[k].L2));
      const dim3 rgrid((maxL + DD_PACK_ROWS - 1) / DD_PACK_ROWS, nprob);
      dd_pack_head<<<nprob, DD_PACK_THREADS, 0, p->stream>>>(c_pp_);
      dd_pack_entries<<<rgrid, DD_PACK_THREADS, 0, p->stream>>>(c_pp_);
      dd_pack_rows<<<rgrid, DD_PACK_THREADS, 0, p->stream>>>(c_pp_);
    }
    PACK_TRY(cudaGetLastError());
    PACK_TRY(cudaStreamSynchronize(p->stream));       // pp / hdr go out of scope; the caller may reuse the matrices
    if (const char* e = getenv("DAFS_DD_PACK_VERIFY")) {
      if (atoi(e) > 0) {
        // test hook: the device-built blocks must equal the host packer's, byte for byte
        std::vector<char> dev(cst);
        PACK_TRY(cudaMemcpy(dev.data(), p->d_cst, cst, cudaMemcpyDeviceToHost));
        for (int k = 0; k < nprob; ++k) {
          const dafs_dd_problem& q = prob[k];
          std::vector<float> hx((size_t)q.L1 * q.L1), hy((size_t)q.L2 * q.L2), hz((size_t)q.L1 * q.L2);
          PACK_TRY(cudaMemcpy(hx.data(), q.p_x, sizeof(float) * hx.size(), cudaMemcpyDeviceToHost));
          PACK_TRY(cudaMemcpy(hy.data(), q.p_y, sizeof(float) * hy.size(), cudaMemcpyDeviceToHost));
          PACK_TRY(cudaMemcpy(hz.data(), q.p_z, sizeof(float) * hz.size(), cudaMemcpyDeviceToHost));
          dafs_dd_problem hq = q;
          hq.p_x = hx.data(); hq.p_y = hy.data(); hq.p_z = hz.data();
          ProbHost ref;
          build_problem(hq, p->want_trace, ref);
          const ProbHost& H = p->host[k];
          if (ref.cst.size() != H.cst_bytes_dev || ref.res_bytes != H.res_bytes)
            return bail_pack(fail(DAFS_CUDA_ECUDA, "pack verify: problem %d: block %zu bytes on the device, %zu on the host", k, H.cst_bytes_dev, ref.cst.size()));
          for (size_t c = 0; c < ref.cst.size(); ++c)
            if (ref.cst[c] != dev[H.cst_base + c])
              return bail_pack(fail(DAFS_CUDA_ECUDA, "pack verify: problem %d: block byte %zu differs (device %d, host %d; stage %d, bytes %zu)", k, c,
                                    (int)(unsigned char)dev[H.cst_base + c], (int)(unsigned char)ref.cst[c], H.stage, ref.cst.size()));
        }
      }
    }
    pack_free();
  } else {
    // one pinned staging buffer -> one H2D copy
    char* stage = (char*)pinned_alloc(std::max<size_t>(cst, 256));
    if (!stage) return bail(fail(DAFS_CUDA_ENOMEM, "pinned staging of %zu bytes failed", cst));
    for (auto& H : p->host) { std::memcpy(stage + H.cst_base, H.cst.data(), H.cst.size()); std::vector<char>().swap(H.cst); }
    cudaError_t e = cudaMemcpyAsync(p->d_cst, stage, cst, cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
    pinned_free(stage);
    PLAN_TRY(e);
  }
#undef PACK_TRY
  std::vector<DDProbRef> dp(nprob);
  std::vector<int32_t> order(nprob);
  for (int k = 0; k < nprob; ++k) {
    const ProbHost& H = p->host[k];
    char* r = p->d_res + H.res_base;
    DDProbRef& D = dp[k];
    D.cst = p->d_cst + H.cst_base;
    D.x_out = reinterpret_cast<uint32_t*>(r + H.sx.out);
    D.y_out = reinterpret_cast<uint32_t*>(r + H.sy.out);
    D.z_out = reinterpret_cast<uint32_t*>(r + H.z_out);
    D.result = reinterpret_cast<float*>(r + H.result);
    D.trace = p->want_trace ? reinterpret_cast<float*>(r + H.trace) : nullptr;
    order[k] = k;
  }
  // Longest expected job first: the batch ends with its slowest problem, and a problem
  // runs the longer the more base-pair constraints it has to reconcile (none: one
  // iteration; many: it usually runs into the -m cap).
  auto cost = [&](const ProbHost& H) {
    const double cells = (double)H.L1 * H.L1 + (double)H.L2 * H.L2 + (double)H.L1 * H.L2;
    return cells * (double)std::min<int64_t>(H.t_max, 1 + (int64_t)H.ncbp);
  };
  std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return cost(p->host[a]) > cost(p->host[b]); });
  PLAN_TRY(cudaMemcpy(p->d_prob, dp.data(), sizeof(DDProbRef) * nprob, cudaMemcpyHostToDevice));
  PLAN_TRY(cudaMemcpy(p->d_order, order.data(), sizeof(int32_t) * nprob, cudaMemcpyHostToDevice));

  // geometry
  dafs::DevInfo prop;
  PLAN_TRY(dafs::dev_info(p->device, &prop));
  const size_t budget = (size_t)prop.smem_optin - 2048;
  size_t ws_per = 0;
  plan_layout(p->lay, m, budget, 1, &p->smem, &ws_per);
  p->team = 1;
  // Fewer problems than SMs (the upper waves of a guide tree): only the latency of one problem
  // counts, so give each Nussinov structure a team of warps -- the largest team that does not
  // push a scratch array out of shared memory.
  {
    // (also when the batch is larger than the SM count but every problem still gets a CTA of its own with the
    // team's shared memory and threads: small problems, several CTAs per SM)
    const bool forced = getenv("DAFS_DD_TEAM") != nullptr;
    int want = 3;
    if (forced) want = std::max(1, std::min(3, atoi(getenv("DAFS_DD_TEAM"))));
    for (int t = want; t > 1; --t) {
      DDLayout lay2 = p->lay; int smem2 = 0; size_t ws2 = 0;
      plan_layout(lay2, m, budget, t, &smem2, &ws2);
      if (!forced) {
        if (lay2.in_smem != p->lay.in_smem) continue;
        if (nprob > prop.sms) {                      // does every problem get its own CTA?
          const bool all2 = (lay2.in_smem == (1u << DD_NITEMS) - 1u);
          const void* fn2 = dd_kernel_fn(all2, t);
          int b2 = 0;
          if (dd_allow_smem(fn2, all2, t, p->device, prop.smem_optin) != cudaSuccess ||
              cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, fn2, 32 * (2 * t + 1), smem2) != cudaSuccess ||
              (int64_t)nprob > (int64_t)prop.sms * b2) { (void)cudaGetLastError(); continue; }
        }
      }
      p->lay = lay2; p->smem = smem2; ws_per = ws2; p->team = t;
      break;
    }
  }
  p->all_smem = (p->lay.in_smem == (1u << DD_NITEMS) - 1u);
  int blocks = 0;
  {
    const void* fn = dd_kernel_fn(p->all_smem, p->team);
    PLAN_TRY(dd_allow_smem(fn, p->all_smem, p->team, p->device, prop.smem_optin));
    PLAN_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, fn, 32 * (2 * p->team + 1), p->smem));
  }
  if (blocks < 1) return bail(fail(DAFS_CUDA_ECUDA, "dd_kernel does not fit on an SM (smem %d)", p->smem));
  p->ctas_per_sm = blocks;
  p->grid = std::max(1, std::min(nprob, prop.sms * blocks));
  p->lay.ws_stride = ws_per;
  PLAN_TRY(dev_malloc(&p->d_ws, ws_per * p->grid));
  p->lay.ws = p->d_ws;
#undef PLAN_TRY
  *out = p;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_dd_plan_create_device(int nprob, const dafs_dd_problem* prob, int want_trace,
                                               dafs_dd_plan** out) {
  return dd_plan_create_impl(nprob, prob, want_trace, true, out);
}

extern "C" int dafs_cuda_dd_plan_create(int nprob, const dafs_dd_problem* prob, int want_trace,
                                        dafs_dd_plan** out) {
  const char* e = getenv("DAFS_DD_DEVICE_PACK");
  if (!(e && atoi(e) > 0) || nprob <= 0 || !prob || !out) return dd_plan_create_impl(nprob, prob, want_trace, false, out);
  // test hook: the same call with the set-up on the device -- the matrices are uploaded as they are and
  // the constant blocks built by the pack kernels
  using namespace dafs;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return dd_plan_create_impl(nprob, prob, want_trace, false, out); }
  size_t tot = 0;
  for (int k = 0; k < nprob; ++k) {
    const dafs_dd_problem& q = prob[k];
    if (q.L1 < 1 || q.L2 < 1 || q.L1 > kMaxL || q.L2 > kMaxL || !q.p_x || !q.p_y || !q.p_z) return dd_plan_create_impl(nprob, prob, want_trace, false, out);
    tot += (size_t)q.L1 * q.L1 + (size_t)q.L2 * q.L2 + (size_t)q.L1 * q.L2;
  }
  float* d = nullptr;
  DAFS_CUDA_TRY(dev_malloc(&d, sizeof(float) * std::max<size_t>(tot, 1)));
  std::vector<dafs_dd_problem> dq(prob, prob + nprob);
  size_t off = 0;
  cudaError_t err = cudaSuccess;
  for (int k = 0; k < nprob && err == cudaSuccess; ++k) {
    const dafs_dd_problem& q = prob[k];
    const size_t nx = (size_t)q.L1 * q.L1, ny = (size_t)q.L2 * q.L2, nz = (size_t)q.L1 * q.L2;
    err = cudaMemcpy(d + off, q.p_x, sizeof(float) * nx, cudaMemcpyHostToDevice); dq[k].p_x = d + off; off += nx;
    if (err == cudaSuccess) err = cudaMemcpy(d + off, q.p_y, sizeof(float) * ny, cudaMemcpyHostToDevice);
    dq[k].p_y = d + off; off += ny;
    if (err == cudaSuccess) err = cudaMemcpy(d + off, q.p_z, sizeof(float) * nz, cudaMemcpyHostToDevice);
    dq[k].p_z = d + off; off += nz;
  }
  int rc = err == cudaSuccess ? dd_plan_create_impl(nprob, dq.data(), want_trace, true, out)
                              : fail(DAFS_CUDA_ECUDA, "upload of the dense matrices failed: %s", cudaGetErrorString(err));
  dev_free(d);
  return rc;
}

extern "C" int dafs_cuda_dd_plan_run(dafs_dd_plan* p) {
  using namespace dafs;
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  p->launched = true;
  if (p->nprob == 0) return DAFS_CUDA_OK;
  DDArgs a;
  a.prob = p->d_prob; a.order = p->d_order; a.nprob = p->nprob; a.queue = p->d_queue; a.lay = p->lay;
  DAFS_CUDA_TRY(cudaMemsetAsync(p->d_queue, 0, sizeof(int), p->stream));
  DAFS_CUDA_TRY(cudaMemsetAsync(p->d_res, 0, p->res_bytes, p->stream));      // completion flags of the last run
  DAFS_CUDA_TRY(cudaEventRecord(p->ev0, p->stream));
  {
    const void* fn = dd_kernel_fn(p->all_smem, p->team);
    void* args[] = {(void*)&a};
    DAFS_CUDA_TRY(cudaLaunchKernel(fn, dim3(p->grid), dim3(32 * (2 * p->team + 1)), args, (size_t)p->smem, p->stream));
  }
  DAFS_CUDA_TRY(cudaGetLastError());
  DAFS_CUDA_TRY(cudaEventRecord(p->ev1, p->stream));
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_dd_plan_sync(dafs_dd_plan* p, float* device_ms) {
  if (!p) return fail(DAFS_CUDA_EINVAL, "plan is NULL");
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  DAFS_CUDA_TRY(cudaStreamSynchronize(p->stream));
  if (p->nprob > 0) DAFS_CUDA_TRY(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
  if (device_ms) *device_ms = p->last_ms;
  return DAFS_CUDA_OK;
}

namespace dafs {
// one problem's slice of the result block -> dafs_dd_result
static void unpack_result(const dafs_dd_plan* p, const char* stage, int k, dafs_dd_result& r) {
  const ProbHost& H = p->host[k];
  const char* s = stage + H.res_base;
  if (r.x) std::memcpy(r.x, s + H.sx.out, sizeof(uint32_t) * H.L1);
  if (r.y) std::memcpy(r.y, s + H.sy.out, sizeof(uint32_t) * H.L2);
  if (r.z) std::memcpy(r.z, s + H.z_out, sizeof(uint32_t) * H.L1);
  const float* rf = reinterpret_cast<const float*>(s + H.result);
  r.score = rf[0];
  r.iters = reinterpret_cast<const int*>(rf)[1];
  r.violated = reinterpret_cast<const unsigned*>(rf)[2];
  r.status = reinterpret_cast<const int*>(rf)[3];
  r.ncbp = H.ncbp;
  for (int c = 0; c < 8; ++c) r.cycles[c] = rf[4 + c];
  if (r.trace && p->want_trace)
    std::memcpy(r.trace, s + H.trace, sizeof(float) * 3 * (size_t)std::min(H.t_max, r.iters + 1));
}
}  // namespace dafs

// Non-blocking harvest: the result block is copied to the host on a second stream while the
// kernel runs; a problem whose completion flag is set in one copy is read from the NEXT copy
// (the flag is written after the results, but one copy reads them in no particular order).
extern "C" int dafs_cuda_dd_plan_poll(dafs_dd_plan* p, dafs_dd_result* res, int32_t* done, int32_t* n_new) {
  using namespace dafs;
  if (!p || !done || !n_new || (p->nprob > 0 && !res)) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  *n_new = 0;
  if (!p->launched) return fail(DAFS_CUDA_EINVAL, "plan was not run");
  if (p->nprob == 0) return DAFS_CUDA_OK;
  DAFS_CUDA_TRY(cudaSetDevice(p->device));
  if (!p->poll_stream) DAFS_CUDA_TRY(cudaStreamCreateWithFlags(&p->poll_stream, cudaStreamNonBlocking));
  if (!p->poll_stage) {
    p->poll_stage = (char*)pinned_alloc(std::max<size_t>(p->res_bytes, 256));
    if (!p->poll_stage) return fail(DAFS_CUDA_ENOMEM, "pinned staging of %zu bytes failed", p->res_bytes);
  }
  // the flags of an earlier run are cleared on the plan's stream just before ev0; a copy on the
  // poll stream must not overtake that
  DAFS_CUDA_TRY(cudaStreamWaitEvent(p->poll_stream, p->ev0, 0));
  std::vector<int> fresh;
  for (int pass = 0; pass < 2; ++pass) {
    DAFS_CUDA_TRY(cudaMemcpyAsync(p->poll_stage, p->d_res, p->res_bytes, cudaMemcpyDeviceToHost, p->poll_stream));
    DAFS_CUDA_TRY(cudaStreamSynchronize(p->poll_stream));
    if (pass == 0) {
      for (int k = 0; k < p->nprob; ++k)
        if (!done[k] && reinterpret_cast<const int*>(p->poll_stage + p->host[k].res_base + p->host[k].result)[12] == 1) fresh.push_back(k);
      if (fresh.empty()) return DAFS_CUDA_OK;
    }
  }
  for (int k : fresh) { unpack_result(p, p->poll_stage, k, res[k]); done[k] = 1; }
  *n_new = (int32_t)fresh.size();
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_dd_plan_fetch(dafs_dd_plan* p, dafs_dd_result* res) {
  using namespace dafs;
  if (!p || (p->nprob > 0 && !res)) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  const int rc = dafs_cuda_dd_plan_sync(p, nullptr);
  if (rc) return rc;
  if (p->nprob == 0) return DAFS_CUDA_OK;
  // x, y, z, result and trace of all problems are one contiguous device block.
  char* stage = (char*)pinned_alloc(std::max<size_t>(p->res_bytes, 256));
  if (!stage) return fail(DAFS_CUDA_ENOMEM, "pinned staging of %zu bytes failed", p->res_bytes);
  cudaError_t e = cudaMemcpyAsync(stage, p->d_res, p->res_bytes, cudaMemcpyDeviceToHost, p->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
  if (e != cudaSuccess) { pinned_free(stage); return fail(DAFS_CUDA_ECUDA, "result copy failed: %s", cudaGetErrorString(e)); }
  for (int k = 0; k < p->nprob; ++k) unpack_result(p, stage, k, res[k]);
  pinned_free(stage);
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_dd_plan_info(dafs_dd_plan* p, dafs_dd_info* info) {
  if (!p || !info) return fail(DAFS_CUDA_EINVAL, "NULL argument");
  info->grid = p->grid; info->block = 32 * (2 * p->team + 1); info->cols_per_lane = p->lay.ts; info->smem_bytes = p->smem;
  info->ctas_per_sm = p->ctas_per_sm; info->items_in_smem = p->lay.in_smem;
  info->workspace_bytes = (int64_t)p->lay.ws_stride * p->grid;
  info->arena_bytes = (int64_t)(p->cst_bytes + p->res_bytes);
  info->total_cbp = p->total_cbp; info->total_entries = p->total_entries; info->last_ms = p->last_ms;
  info->packed_on_device = p->packed_on_device ? 1 : 0;
  return DAFS_CUDA_OK;
}

extern "C" void dafs_cuda_dd_plan_destroy(dafs_dd_plan* p) {
  if (!p) return;
  cudaSetDevice(p->device);
  if (p->stream) cudaStreamSynchronize(p->stream);
  dafs::free_dd_plan(p);
  delete p;
}

extern "C" int dafs_cuda_dd_solve_batch(int nprob, const dafs_dd_problem* prob, dafs_dd_result* res) {
  dafs_dd_plan* p = nullptr;
  int rc = dafs_cuda_dd_plan_create(nprob, prob, 0, &p);
  if (rc) return rc;
  rc = dafs_cuda_dd_plan_run(p);
  if (!rc) rc = dafs_cuda_dd_plan_fetch(p, res);
  dafs_cuda_dd_plan_destroy(p);
  return rc;
}

// ---- single decodes ------------------------------------------------------------
namespace dafs {

__global__ void __launch_bounds__(32) nussinov_single_kernel(const char* blk, float* q, const NusScratch sc, float th,
                                                             uint32_t* out, float* score) {
  const DDNus S = dd_make_nus(blk, *reinterpret_cast<const DDNusHdr*>(blk), q);
  for (int e = threadIdx.x; e < S.nd; e += 32) sc.es[e] = -INFINITY;
  __syncwarp();
  const float v = nussinov_warp<1>(S, sc, th, threadIdx.x, 0, 1, nullptr);
  if (threadIdx.x == 0) *score = v;
  for (int i = threadIdx.x; i < S.L; i += 32) out[i] = sc.ss[i];
}

__global__ void __launch_bounds__(32) nw_single_kernel(const DDAln A, const NwScratch sc, uint32_t* al, float* score,
                                                       int* bad) {
  const float v = nw_warp(A, sc, threadIdx.x, bad, nullptr);
  if (threadIdx.x == 0) *score = v;
  for (int i = threadIdx.x; i < A.L1; i += 32) al[i] = sc.zres[i];
}

struct DevBuf {   // one cudaMalloc, carved
  char* base = nullptr; size_t size = 0, used = 0;
  ~DevBuf() { dev_free(base); }
  template <class T> T* take(size_t n) { used = up16(used); T* p = reinterpret_cast<T*>(base + used); used += sizeof(T) * std::max<size_t>(n, 1); return p; }
};

}  // namespace dafs

extern "C" int dafs_cuda_nussinov_decode(int L, float th, float w, const float* p, const float* q,
                                         uint32_t* ss, float* score) {
  using namespace dafs;
  if (L < 1 || L > kMaxL || !p || !ss) return fail(DAFS_CUDA_EINVAL, "bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)"); }
  if (!q) w = 1.0f;   // plain overload: s = p - th == 1*(p - th) - 0 exactly
  // candidates: s > 0, span > 2 (nussinov.cpp:234-237); nothing else is ever read
  std::vector<Ent> ents;
  std::vector<float> qv;
  for (int i = 0; i < L; ++i)
    for (int j = i + 3; j < L; ++j) {
      const float pv = p[(size_t)i * L + j], qq = q ? q[(size_t)i * L + j] : 0.0f;
      const float s = w * (pv - th) - qq;
      if (s > 0.0) ents.push_back({i, j, pv, 0, true});
    }
  NusHost H;
  std::vector<char> blk(sizeof(DDNusHdr), 0);
  pack_nus(H, blk, L, w, ents);          // sorts ents into column order
  std::memcpy(blk.data(), &H.h, sizeof(DDNusHdr));
  blk.resize(up16(blk.size()));
  const int n = H.h.n, nI = std::max(H.h.nI, 1), nJ = std::max(H.h.nJ, 1);
  qv.resize(std::max(n, 1));
  for (int e = 0; e < n; ++e) qv[e] = q ? q[(size_t)ents[e].i * L + ents[e].j] : 0.0f;
  const bool tr16 = H.max_col > 120;
  const int lp = (nJ + 2) & ~1;
  const size_t ntr = (size_t)nI * nJ + 1;
  DevBuf B;
  B.size = blk.size() + 64 * 16 + 4 * (size_t)(2 * (size_t)std::max(n, 1) + 33 * (size_t)lp + 2 * L + L / 2 + 16) + (tr16 ? 2 : 1) * ntr + 64;
  DAFS_CUDA_TRY(dev_malloc(&B.base, B.size));
  NusScratch sc;
  char* d_blk = B.take<char>(blk.size());
  float* d_q = B.take<float>(n);
  sc.es = B.take<float>(n); sc.row = B.take<float>(32 * (size_t)lp); sc.brow = B.take<float>(lp);
  sc.prog = B.take<int>(4); sc.score_slot = B.take<float>(1);
  sc.ss = B.take<uint32_t>(L); sc.stk = B.take<uint32_t>(L / 2 + 4);
  uint32_t* d_out = B.take<uint32_t>(L); float* d_score = B.take<float>(1);
  sc.tr = tr16 ? (void*)B.take<uint16_t>(ntr) : (void*)B.take<uint8_t>(ntr);
  sc.lp = lp; sc.trs = nJ; sc.tr16 = tr16;
  DAFS_CUDA_TRY(cudaMemcpy(d_blk, blk.data(), blk.size(), cudaMemcpyHostToDevice));
  DAFS_CUDA_TRY(cudaMemcpy(d_q, qv.data(), sizeof(float) * std::max(n, 1), cudaMemcpyHostToDevice));
  nussinov_single_kernel<<<1, 32>>>(d_blk, d_q, sc, th, d_out, d_score);
  DAFS_CUDA_TRY(cudaGetLastError());
  DAFS_CUDA_TRY(cudaMemcpy(ss, d_out, sizeof(uint32_t) * L, cudaMemcpyDeviceToHost));
  float sv = 0;
  DAFS_CUDA_TRY(cudaMemcpy(&sv, d_score, sizeof(float), cudaMemcpyDeviceToHost));
  if (score) *score = sv;
  return DAFS_CUDA_OK;
}

extern "C" int dafs_cuda_nw_decode(int L1, int L2, float th, const float* p, const float* q, uint32_t* al,
                                   float* score) {
  using namespace dafs;
  if (L1 < 1 || L2 < 1 || L1 > kMaxL || L2 > kMaxL || !p || !al) return fail(DAFS_CUDA_EINVAL, "bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)"); }
  std::vector<uint2> env;
  build_envelope(L1, L2, th, p, env);
  // dense bands of p and q over the envelope (cells outside it are never evaluated, :262-264)
  std::vector<int> boff(L1 + 2, 0);
  std::vector<float> pv, qv;
  int bw = 1;
  for (int r = 1; r <= L1; ++r) {
    boff[r] = (int)pv.size();
    const int ef = std::max(1, (int)env[r].x), el = (int)env[r].y;
    bw = std::max(bw, el - ef + 1);
    for (int k = ef; k <= el; ++k) {
      pv.push_back(p[(size_t)(r - 1) * L2 + (k - 1)]);
      qv.push_back(q ? q[(size_t)(r - 1) * L2 + (k - 1)] : 0.0f);
    }
  }
  boff[L1 + 1] = (int)pv.size();
  const int n = (int)pv.size(), ts = (bw + 31) / 32;
  DevBuf B;
  B.size = 16 * 16 + 4 * ((size_t)(L1 + 2) * (3 + 2 * ts) + 2 * (size_t)std::max(n, 1) + 2 * (L2 + 2) + 2 * (size_t)L1 + 8) + 64;
  DAFS_CUDA_TRY(dev_malloc(&B.base, B.size));
  DDAln A{}; NwScratch sc{};
  int* d_boff = B.take<int>(L1 + 2);
  float* d_p = B.take<float>(n); float* d_q = B.take<float>(n);
  uint2* d_env = B.take<uint2>(L1 + 1);
  sc.rowbuf = B.take<float>(2 * (L2 + 2)); sc.trz = B.take<uint2>((size_t)(L1 + 1) * ts); sc.zres = B.take<uint32_t>(L1);
  sc.ts = ts;
  { const char* e = getenv("DAFS_DD_NW_CHUNKED"); sc.chunked = (e && atoi(e) > 0) ? 1 : 0; }
  uint32_t* d_al = B.take<uint32_t>(L1); float* d_score = B.take<float>(1); int* d_bad = B.take<int>(1);
  DAFS_CUDA_TRY(cudaMemcpy(d_boff, boff.data(), sizeof(int) * (L1 + 2), cudaMemcpyHostToDevice));
  if (n) {
    DAFS_CUDA_TRY(cudaMemcpy(d_p, pv.data(), sizeof(float) * n, cudaMemcpyHostToDevice));
    DAFS_CUDA_TRY(cudaMemcpy(d_q, qv.data(), sizeof(float) * n, cudaMemcpyHostToDevice));
  }
  DAFS_CUDA_TRY(cudaMemcpy(d_env, env.data(), sizeof(uint2) * (L1 + 1), cudaMemcpyHostToDevice));
  DAFS_CUDA_TRY(cudaMemset(d_bad, 0, sizeof(int)));
  DAFS_CUDA_TRY(cudaMemset(d_score, 0, sizeof(float)));
  A.L1 = L1; A.L2 = L2; A.th = th; A.env = d_env; A.boff = d_boff; A.pb = d_p; A.qb = d_q;
  nw_single_kernel<<<1, 32>>>(A, sc, d_al, d_score, d_bad);
  DAFS_CUDA_TRY(cudaGetLastError());
  DAFS_CUDA_TRY(cudaMemcpy(al, d_al, sizeof(uint32_t) * L1, cudaMemcpyDeviceToHost));
  float sv = 0; int bad = 0;
  DAFS_CUDA_TRY(cudaMemcpy(&sv, d_score, sizeof(float), cudaMemcpyDeviceToHost));
  DAFS_CUDA_TRY(cudaMemcpy(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost));
  if (score) *score = sv;
  if (bad) return fail(DAFS_CUDA_ECUDA, "NW traceback met an unfilled cell (the reference does not terminate on this input)");
  return DAFS_CUDA_OK;
}

// ---- allocation-free single NW decode --------------------------------------------
// dafs_cuda_nw_decode_static: same arguments and results as dafs_cuda_nw_decode, for
// callers that may not call cudaMalloc (code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  Every device byte it needs -- the band image of
// p / q, the envelope, the row buffers, the traceback planes, the result -- lives in
// ONE file-scope device array of DAFS_CUDA_NW_STATIC_BYTES; the host image of the
// inputs is a file-scope host array of the same size.  No cudaMalloc / cudaFree /
// cudaHostAlloc, no host heap, no allocating library routine on this path.
//   * The array is shared by every call (and held by every device that loads the
//     library until exit); it keeps the last call's contents.  What must start at zero
//     (score, the bad-cell flag) is part of the image the call uploads; everything
//     else is written before it is read.
//   * Calls are serialised by a mutex on the host and run on the legacy default stream;
//     two calls that overlap on the GPU (other processes share nothing) would race.
namespace dafs {

static __device__ __align__(256) unsigned char g_nw_static[DAFS_CUDA_NW_STATIC_BYTES];
static unsigned char g_nw_static_host[DAFS_CUDA_NW_STATIC_BYTES];   // host image of the inputs
static std::mutex g_nw_static_mu;
static unsigned char* g_nw_static_addr[64] = {};                    // device address, per device

// `g_nw_static` in host code is a host-side stand-in, not a device address: ask the
// runtime once per device (with that device current) for the real one.
static int nw_static_base(unsigned char** out) {
  int dev = 0;
  DAFS_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(DAFS_CUDA_ENODEV, "device index %d outside [0, 64)", dev);
  if (!g_nw_static_addr[dev]) {
    void* p = nullptr;
    DAFS_CUDA_TRY(cudaGetSymbolAddress(&p, g_nw_static));
    if (!p) return fail(DAFS_CUDA_ECUDA, "cudaGetSymbolAddress returned NULL");
    g_nw_static_addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_nw_static_addr[dev];
  return DAFS_CUDA_OK;
}

// SparseNeedlemanWunsch::initialize, src/needleman_wunsch.cpp:198-253, into env[L1+1]
static void envelope_into(int L1, int L2, float th, const float* p, uint2* env) {
  for (int i = 0; i <= L1; ++i) env[i] = make_uint2(0, 0);
  for (int i = 1; i <= L1; ++i) {
    const float* row = p + (size_t)(i - 1) * L2;
    int kf = 0, kl = 0;
    for (int k = 1; k <= L2 && !kf; ++k) if (row[k - 1] - th >= 0.0) kf = k;
    if (!kf) { env[i] = env[i - 1]; continue; }
    for (int k = L2; k >= 1 && !kl; --k) if (row[k - 1] - th >= 0.0) kl = k;
    env[i - 1].x = std::min(env[i - 1].x, (unsigned)(kf - 1)); env[i].x = (unsigned)kf;
    env[i - 1].y = std::max(env[i - 1].y, (unsigned)(kl - 1)); env[i].y = (unsigned)kl;
  }
  env[L1].y = (unsigned)L2;
  unsigned lo = (unsigned)L2, hi = 0;
  for (int i = L1; i >= 1; --i) env[i].x = lo = std::min(lo, env[i].x);
  for (int i = 0; i <= L1; ++i) env[i].y = hi = std::max(hi, env[i].y);
  for (int i = 1; i <= L1; ++i) if (env[i - 1].y < env[i].x) env[i].x = env[i - 1].y;
}

}  // namespace dafs

extern "C" int dafs_cuda_nw_decode_static(int L1, int L2, float th, const float* p, const float* q,
                                          uint32_t* al, float* score) {
  using namespace dafs;
  if (L1 < 1 || L2 < 1 || !p || !al) return fail(DAFS_CUDA_EINVAL, "bad argument");
  if (L1 > DAFS_CUDA_NW_STATIC_MAXL || L2 > DAFS_CUDA_NW_STATIC_MAXL)
    return fail(DAFS_CUDA_EINVAL, "dafs_cuda_nw_decode_static: L1=%d L2=%d beyond the static limit of %d (use dafs_cuda_nw_decode)",
                L1, L2, DAFS_CUDA_NW_STATIC_MAXL);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(DAFS_CUDA_ENODEV, "no CUDA device visible (libdafs_cuda has no CPU fallback)"); }
  std::lock_guard<std::mutex> lock(g_nw_static_mu);
  unsigned char* base = nullptr;
  if (const int rc = nw_static_base(&base)) return rc;

  // image = {score, bad, pad} | al[L1] | env[L1+1] | boff[L1+2] | p band | q band ; then scratch
  unsigned char* h = g_nw_static_host;
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t o = off; off = up16(off + std::max<size_t>(bytes, 16)); return o; };
  const size_t o_hdr = take(16), o_al = take(sizeof(uint32_t) * L1), o_env = take(sizeof(uint2) * (L1 + 1)),
               o_boff = take(sizeof(int) * (L1 + 2));
  std::memset(h + o_hdr, 0, 16);
  uint2* env = reinterpret_cast<uint2*>(h + o_env);
  int* boff = reinterpret_cast<int*>(h + o_boff);
  envelope_into(L1, L2, th, p, env);
  size_t n = 0; int bw = 1;
  boff[0] = 0;
  for (int r = 1; r <= L1; ++r) {
    boff[r] = (int)n;
    const int ef = std::max(1, (int)env[r].x), el = (int)env[r].y;
    if (el >= ef) { n += (size_t)(el - ef + 1); bw = std::max(bw, el - ef + 1); }
  }
  boff[L1 + 1] = (int)n;
  const int ts = (bw + 31) / 32;
  const size_t o_p = take(sizeof(float) * n), o_q = take(sizeof(float) * n);
  const size_t image = off;
  const size_t o_row = take(sizeof(float) * 2 * (size_t)(L2 + 2)), o_trz = take(sizeof(uint2) * (size_t)(L1 + 1) * ts),
               o_zres = take(sizeof(uint32_t) * L1);
  if (off > (size_t)DAFS_CUDA_NW_STATIC_BYTES)
    return fail(DAFS_CUDA_EINVAL, "dafs_cuda_nw_decode_static: %d x %d with %zu envelope cells needs %zu bytes, the static scratch holds %d (use dafs_cuda_nw_decode)",
                L1, L2, n, off, DAFS_CUDA_NW_STATIC_BYTES);
  float* pb = reinterpret_cast<float*>(h + o_p);
  float* qb = reinterpret_cast<float*>(h + o_q);
  for (int r = 1; r <= L1; ++r) {
    const int ef = std::max(1, (int)env[r].x), el = (int)env[r].y;
    size_t w = (size_t)boff[r];
    for (int k = ef; k <= el; ++k, ++w) {
      pb[w] = p[(size_t)(r - 1) * L2 + (k - 1)];
      qb[w] = q ? q[(size_t)(r - 1) * L2 + (k - 1)] : 0.0f;
    }
  }
  DAFS_CUDA_TRY(cudaMemcpy(base, h, image, cudaMemcpyHostToDevice));      // also zeroes score / bad
  DDAln A{}; NwScratch sc{};
  A.L1 = L1; A.L2 = L2; A.th = th;
  A.env = reinterpret_cast<const uint2*>(base + o_env); A.boff = reinterpret_cast<const int*>(base + o_boff);
  A.pb = reinterpret_cast<const float*>(base + o_p); A.qb = reinterpret_cast<const float*>(base + o_q);
  sc.rowbuf = reinterpret_cast<float*>(base + o_row); sc.trz = reinterpret_cast<uint2*>(base + o_trz);
  sc.zres = reinterpret_cast<uint32_t*>(base + o_zres); sc.ts = ts;
  { const char* e = getenv("DAFS_DD_NW_CHUNKED"); sc.chunked = (e && atoi(e) > 0) ? 1 : 0; }
  nw_single_kernel<<<1, 32>>>(A, sc, reinterpret_cast<uint32_t*>(base + o_al), reinterpret_cast<float*>(base + o_hdr),
                              reinterpret_cast<int*>(base + o_hdr + 4));
  DAFS_CUDA_TRY(cudaGetLastError());
  DAFS_CUDA_TRY(cudaMemcpy(h, base, o_env, cudaMemcpyDeviceToHost));      // {score, bad} | al
  std::memcpy(al, h + o_al, sizeof(uint32_t) * L1);
  if (score) std::memcpy(score, h + o_hdr, sizeof(float));
  int bad = 0; std::memcpy(&bad, h + o_hdr + 4, sizeof(int));
  if (bad) return fail(DAFS_CUDA_ECUDA, "NW traceback met an unfilled cell (the reference does not terminate on this input)");
  return DAFS_CUDA_OK;
}
