/* TEST INFRASTRUCTURE ONLY -- never linked into, or called from, the product.
 *
 * Plain-C restatement of the reference's progressive stage: the guide tree and the
 * recursion that aligns two sub-alignments per tree node (DAFS::run,
 * src/dafs.cpp:1829-1838):
 *   build_tree                          src/dafs.cpp:447-495
 *   average_matching_probability        src/dafs.cpp:513-555
 *   average_basepairing_probability     src/dafs.cpp:557-607 (use_alifold = false)
 *   project_alignment                   src/dafs.cpp:766-826
 *   project_secondary_structure         src/dafs.cpp:828-876
 *   align_alignments(ss, aln, a1, a2)   src/dafs.cpp:913-951 (use_bp_update_ = false)
 *   align(ss, aln, ch)                  src/dafs.cpp:1518-1537
 * The solver it calls per node is oracle_dd_solve (dd_oracle.c).  Pinned bit-for-bit
 * against the reference's own functions compiled from /root/reference
 * (oracle/_ref: ref_progressive, tests/test_oracle_progressive.py).
 *
 * Data layout = dafs_csr_batch: mp holds all i<j pairs row-major (rows = len[i]),
 * bp one unit per sequence; row_ptr blocks have rows+1 entries relative to nz_off. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CUTOFF 0.01 /* src/dafs.cpp:65, a double */

float oracle_dd_solve(int L1, int L2, int N1, int N2, const float* p_x, const float* p_y,
                      const float* p_z, float w_, float th_s, float min_th_s, float th_a,
                      float eta0, int t_max, uint32_t* x, uint32_t* y, uint32_t* z, int* iters,
                      uint32_t* violated_out, int64_t* ncbp_out, float* trace, int trace_cap);

/* ---- build_tree, src/dafs.cpp:447-495 ------------------------------------------
 * A max-queue of (similarity, (i, j)) entries, ordered like std::pair (value, then i,
 * then j): the two live clusters of the best entry are merged into node n, whose
 * distance to every other live cluster ii is (d[ii][l] + d[ii][r]) * value / 2 (float),
 * stored in slot l.  idx[] maps a queue label to the matrix slot of its cluster. */
typedef struct { float v; uint32_t a, b; } qent_t;
static int q_less(const qent_t* x, const qent_t* y) { /* std::pair operator< */
  if (x->v < y->v) return 1;
  if (y->v < x->v) return 0;
  if (x->a != y->a) return x->a < y->a;
  return x->b < y->b;
}
static void q_push(qent_t** h, size_t* n, size_t* cap, qent_t e) {
  if (*n == *cap) { *cap = *cap ? 2 * *cap : 64; *h = (qent_t*)realloc(*h, sizeof(qent_t) * *cap); }
  size_t k = (*n)++;
  (*h)[k] = e;
  while (k > 0) {
    const size_t p = (k - 1) / 2;
    if (!q_less(&(*h)[p], &(*h)[k])) break;
    const qent_t t = (*h)[p]; (*h)[p] = (*h)[k]; (*h)[k] = t;
    k = p;
  }
}
static qent_t q_pop(qent_t* h, size_t* n) {
  const qent_t top = h[0];
  h[0] = h[--(*n)];
  size_t k = 0;
  for (;;) {
    size_t c = 2 * k + 1;
    if (c >= *n) break;
    if (c + 1 < *n && q_less(&h[c], &h[c + 1])) ++c;
    if (!q_less(&h[k], &h[c])) break;
    const qent_t t = h[k]; h[k] = h[c]; h[c] = t;
    k = c;
  }
  return top;
}

void oracle_build_tree(int n0, const float* sim, int* left, int* right, float* score) {
  const uint32_t N = (uint32_t)n0, none = 0xffffffffu;
  uint32_t n = N;
  for (uint32_t k = 0; k < 2 * N - 1; ++k) { left[k] = -1; right[k] = -1; score[k] = 0.0f; }
  float* d = (float*)calloc((size_t)N * N, sizeof(float));
  uint32_t* idx = (uint32_t*)malloc(sizeof(uint32_t) * (2 * N - 1));
  for (uint32_t i = 0; i < 2 * N - 1; ++i) idx[i] = i < N ? i : none;
  qent_t* h = NULL; size_t hn = 0, hc = 0;
  for (uint32_t i = 0; i + 1 < N; ++i)
    for (uint32_t j = i + 1; j < N; ++j) {
      d[i * N + j] = d[j * N + i] = sim[i * N + j];
      const qent_t e = {sim[i * N + j], i, j};
      q_push(&h, &hn, &hc, e);
    }
  while (hn) {
    const qent_t t = q_pop(h, &hn);
    if (idx[t.a] != none && idx[t.b] != none) {
      const uint32_t l = idx[t.a], r = idx[t.b];
      idx[t.a] = idx[t.b] = none;
      for (uint32_t i = 0; i < n; ++i)
        if (idx[i] != none) {
          const uint32_t ii = idx[i];
          const float v = (d[ii * N + l] + d[ii * N + r]) * t.v / 2;
          d[ii * N + l] = d[l * N + ii] = v;
          const qent_t e = {v, i, n};
          q_push(&h, &hn, &hc, e);
        }
      left[n] = (int)t.a; right[n] = (int)t.b; score[n] = t.v;
      idx[n++] = l;
    }
  }
  free(h); free(idx); free(d);
}

/* ---- alignments ------------------------------------------------------------------
 * ALN = rows of (sequence index, gap pattern over the alignment's columns). */
typedef struct { int nrows, L; int* seq; uint8_t* bits; /* [nrows * L] */ } aln_t;
static void aln_free(aln_t* a) { free(a->seq); free(a->bits); memset(a, 0, sizeof *a); }

typedef struct { const int* row_ptr; const uint32_t* col; const float* val; } view_t;

typedef struct {
  int n; const int* len;
  view_t* mp;   /* [n*n]: i<j the caller's, i>j transposes (owned), i==i unused here */
  view_t* bp;   /* [n] */
  float w, th_s, th_a, eta0; int t_max;
  int* log; float* log_score; int nlog;
} ctx_t;

/* average_matching_probability, src/dafs.cpp:513-555 */
static float* average_mp(const ctx_t* c, const aln_t* a1, const aln_t* a2) {
  const int L1 = a1->L, L2 = a2->L;
  const unsigned N1 = (unsigned)a1->nrows, N2 = (unsigned)a2->nrows;
  float* p = (float*)calloc((size_t)L1 * L2, sizeof(float));
  for (int r1 = 0; r1 < a1->nrows; ++r1)
    for (int r2 = 0; r2 < a2->nrows; ++r2) {
      const view_t* m = &c->mp[(size_t)a1->seq[r1] * c->n + a2->seq[r2]];
      const uint8_t* g1 = a1->bits + (size_t)r1 * L1;
      const uint8_t* g2 = a2->bits + (size_t)r2 * L2;
      for (int i = 0, ii = 0; i != L1; ++i) {
        if (!g1[i]) continue;
        int x = m->row_ptr[ii];
        const int xe = m->row_ptr[ii + 1];
        for (int j = 0, jj = 0; j != L2 && x != xe; ++j) {
          if (!g2[j]) continue;
          if ((uint32_t)jj == m->col[x]) { p[(size_t)i * L2 + j] += m->val[x] / (N1 * N2); ++x; }
          ++jj;
        }
        ++ii;
      }
    }
  for (size_t k = 0; k < (size_t)L1 * L2; ++k) {
    if (p[k] <= CUTOFF) p[k] = 0.0f;
    if (p[k] > 1.0) p[k] = 1.0f;
  }
  return p;
}

/* average_basepairing_probability, src/dafs.cpp:557-607 with use_alifold == false */
static float* average_bp(const ctx_t* c, const aln_t* a) {
  const int L = a->L;
  const unsigned N = (unsigned)a->nrows;
  float* p = (float*)calloc((size_t)L * L, sizeof(float));
  int* idx = (int*)malloc(sizeof(int) * (L > 0 ? L : 1));
  for (int r = 0; r < a->nrows; ++r) {
    const int s = a->seq[r];
    const uint8_t* g = a->bits + (size_t)r * L;
    for (int i = 0, j = 0; i != L; ++i) if (g[i]) idx[j++] = i;
    const view_t* b = &c->bp[s];
    for (int i = 0; i < c->len[s]; ++i)
      for (int k = b->row_ptr[i]; k < b->row_ptr[i + 1]; ++k)
        p[(size_t)idx[i] * L + idx[b->col[k]]] += b->val[k] / N;
  }
  for (int i = 0; i + 1 < L; ++i)
    for (int j = i + 1; j < L; ++j)
      if (p[(size_t)i * L + j] <= CUTOFF) p[(size_t)i * L + j] = 0.0f;
  free(idx);
  return p;
}

/* project_alignment, src/dafs.cpp:766-826: z[i] = column of aln2 matched to column i of
 * aln1 (or -1u); unmatched columns of aln2 are inserted before the next matched one. */
static void project_alignment(aln_t* out, const aln_t* a1, const aln_t* a2, const uint32_t* z) {
  const int L1 = a1->L, L2 = a2->L;
  int cmatch = 0;
  for (int i = 0; i < L1; ++i) if (z[i] != 0xffffffffu) ++cmatch;
  const int L = L1 + L2 - cmatch;
  out->nrows = a1->nrows + a2->nrows; out->L = L;
  out->seq = (int*)malloc(sizeof(int) * out->nrows);
  out->bits = (uint8_t*)calloc((size_t)out->nrows * (L > 0 ? L : 1), 1);
  for (int q = 0; q < a1->nrows; ++q) {
    uint8_t* o = out->bits + (size_t)q * L;
    const uint8_t* g = a1->bits + (size_t)q * L1;
    out->seq[q] = a1->seq[q];
    uint32_t r = 0, k = 0;
    for (int i = 0; i < L1; ++i) {
      if (z[i] != 0xffffffffu) {
        while (k < z[i]) { o[r++] = 0; k++; }
        o[r++] = g[i]; ++k;
      } else o[r++] = g[i];
    }
    while (k < (uint32_t)L2) { o[r++] = 0; k++; }
  }
  for (int q = 0; q < a2->nrows; ++q) {
    uint8_t* o = out->bits + (size_t)(a1->nrows + q) * L;
    const uint8_t* g = a2->bits + (size_t)q * L2;
    out->seq[a1->nrows + q] = a2->seq[q];
    uint32_t r = 0, k = 0;
    for (int i = 0; i < L1; ++i) {
      if (z[i] != 0xffffffffu) {
        while (k < z[i]) o[r++] = g[k++];
        o[r++] = g[k++];
      } else o[r++] = 0;
    }
    while (k < (uint32_t)L2) o[r++] = g[k++];
  }
}

/* project_secondary_structure, src/dafs.cpp:828-876, followed by the consensus of
 * align_alignments, :941-945: ss[i] = xx[i] where xx[i] == yy[i] (also -1u == -1u). */
static void project_ss(uint32_t* ss, int L, const uint32_t* x, int L1, const uint32_t* y, int L2, const uint32_t* z) {
  const uint32_t none = 0xffffffffu;
  uint32_t* idx1 = (uint32_t*)malloc(sizeof(uint32_t) * (L1 > 0 ? L1 : 1));
  uint32_t* idx2 = (uint32_t*)malloc(sizeof(uint32_t) * (L2 > 0 ? L2 : 1));
  uint32_t r = 0, k = 0;
  for (int i = 0; i < L1; ++i) {
    if (z[i] != none) {
      while (k < z[i]) { idx2[k] = r; ++r; ++k; }
      idx1[i] = r; idx2[k] = r; ++r; ++k;
    } else { idx1[i] = r; ++r; }
  }
  while (k < (uint32_t)L2) { idx2[k] = r; ++r; ++k; }
  uint32_t* xx = (uint32_t*)malloc(sizeof(uint32_t) * (L > 0 ? L : 1));
  uint32_t* yy = (uint32_t*)malloc(sizeof(uint32_t) * (L > 0 ? L : 1));
  for (int i = 0; i < L; ++i) xx[i] = yy[i] = none;
  for (int i = 0; i < L1; ++i) if (x[i] != none) xx[idx1[i]] = idx1[x[i]];
  for (int q = 0; q < L2; ++q) if (y[q] != none) yy[idx2[q]] = idx2[y[q]];
  for (int i = 0; i < L; ++i) ss[i] = xx[i] == yy[i] ? xx[i] : none;
  free(idx1); free(idx2); free(xx); free(yy);
}

/* align(ss, aln, ch), src/dafs.cpp:1518-1537, and align_alignments(ss, ...), :913-951 */
static float align_node(ctx_t* c, const int* left, const int* right, int ch, aln_t* out, uint32_t** ss_out) {
  if (left[ch] < 0) {
    out->nrows = 1; out->L = c->len[ch];
    out->seq = (int*)malloc(sizeof(int)); out->seq[0] = ch;
    out->bits = (uint8_t*)malloc(c->len[ch] > 0 ? c->len[ch] : 1);
    memset(out->bits, 1, c->len[ch]);
    return 0.0f;
  }
  aln_t a1 = {0}, a2 = {0};
  align_node(c, left, right, left[ch], &a1, NULL);      /* plain overload, :1499-1516 */
  align_node(c, left, right, right[ch], &a2, NULL);
  float* p_x = average_bp(c, &a1);
  float* p_y = average_bp(c, &a2);
  float* p_z = average_mp(c, &a1, &a2);
  uint32_t* x = (uint32_t*)malloc(sizeof(uint32_t) * a1.L);
  uint32_t* y = (uint32_t*)malloc(sizeof(uint32_t) * a2.L);
  uint32_t* z = (uint32_t*)malloc(sizeof(uint32_t) * a1.L);
  int iters = 0; uint32_t violated = 0; int64_t ncbp = 0;
  const float s = oracle_dd_solve(a1.L, a2.L, a1.nrows, a2.nrows, p_x, p_y, p_z, c->w, c->th_s, c->th_s,
                                  c->th_a, c->eta0, c->t_max, x, y, z, &iters, &violated, &ncbp, NULL, 0);
  if (c->log) {
    int* l = c->log + 4 * c->nlog;
    l[0] = a1.L; l[1] = a2.L; l[2] = iters; l[3] = (int)violated;
    c->log_score[c->nlog] = s;
  }
  c->nlog++;
  project_alignment(out, &a1, &a2, z);
  if (ss_out) {
    *ss_out = (uint32_t*)malloc(sizeof(uint32_t) * (out->L > 0 ? out->L : 1));
    project_ss(*ss_out, out->L, x, a1.L, y, a2.L, z);
  }
  free(p_x); free(p_y); free(p_z); free(x); free(y); free(z);
  aln_free(&a1); aln_free(&a2);
  return s;
}

/* Same contract as ref_progressive (oracle/ref_shim.cpp) with the tree given. */
float oracle_progressive(int n, const int* len, const long long* mp_row_off, const int* mp_row_ptr,
                         const long long* mp_nz_off, const uint32_t* mp_col, const float* mp_val,
                         const long long* bp_row_off, const int* bp_row_ptr, const long long* bp_nz_off,
                         const uint32_t* bp_col, const float* bp_val, const int* left, const int* right,
                         float w, float th_s, float th_a, float eta0, int t_max, int cap, int* L_out,
                         int* seq_order, unsigned char* bits, uint32_t* ss_out, int* solve_log, float* solve_score) {
  ctx_t c;
  memset(&c, 0, sizeof c);
  c.n = n; c.len = len; c.w = w; c.th_s = th_s; c.th_a = th_a; c.eta0 = eta0; c.t_max = t_max;
  c.log = solve_log; c.log_score = solve_score;
  c.mp = (view_t*)calloc((size_t)n * n, sizeof(view_t));
  c.bp = (view_t*)calloc((size_t)n, sizeof(view_t));
  for (int i = 0; i < n; ++i) {
    c.bp[i].row_ptr = bp_row_ptr + bp_row_off[i]; c.bp[i].col = bp_col + bp_nz_off[i]; c.bp[i].val = bp_val + bp_nz_off[i];
  }
  long long u = 0;
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j, ++u) {
      view_t* m = &c.mp[(size_t)i * n + j];
      m->row_ptr = mp_row_ptr + mp_row_off[u]; m->col = mp_col + mp_nz_off[u]; m->val = mp_val + mp_nz_off[u];
      /* transpose_mp, src/dafs.cpp:155-167 (run() :1797-1799): ascending rows inside a column */
      const int nn = m->row_ptr[len[i]];
      int* trp = (int*)calloc((size_t)len[j] + 1, sizeof(int));
      uint32_t* tc = (uint32_t*)malloc(sizeof(uint32_t) * (nn > 0 ? nn : 1));
      float* tv = (float*)malloc(sizeof(float) * (nn > 0 ? nn : 1));
      for (int k = 0; k < nn; ++k) ++trp[m->col[k] + 1];
      for (int q = 0; q < len[j]; ++q) trp[q + 1] += trp[q];
      int* cur = (int*)malloc(sizeof(int) * (len[j] + 1));
      memcpy(cur, trp, sizeof(int) * (len[j] + 1));
      for (int r = 0; r < len[i]; ++r)
        for (int k = m->row_ptr[r]; k < m->row_ptr[r + 1]; ++k) {
          const int pos = cur[m->col[k]]++;
          tc[pos] = (uint32_t)r; tv[pos] = m->val[k];
        }
      free(cur);
      view_t* t = &c.mp[(size_t)j * n + i];
      t->row_ptr = trp; t->col = tc; t->val = tv;
    }
  aln_t root = {0};
  uint32_t* ss = NULL;
  const float s = align_node(&c, left, right, 2 * n - 2, &root, &ss);
  *L_out = root.L;
  if (root.L <= cap) {
    for (int r = 0; r < root.nrows; ++r) {
      seq_order[r] = root.seq[r];
      memcpy(bits + (size_t)r * cap, root.bits + (size_t)r * root.L, root.L);
    }
    if (ss) memcpy(ss_out, ss, sizeof(uint32_t) * root.L);
  }
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < i; ++j) {
      view_t* t = &c.mp[(size_t)i * n + j];
      free((void*)t->row_ptr); free((void*)t->col); free((void*)t->val);
    }
  free(c.mp); free(c.bp); free(ss); aln_free(&root);
  return s;
}
