/* TEST INFRASTRUCTURE ONLY -- never linked into, or called from, the product.
 *
 * Plain-C restatement of the reference's probabilistic consistency transformation,
 * the host stage that consumes the pair-HMM and McCaskill posteriors
 * (DAFS::run, src/dafs.cpp:1796-1827):
 *   transpose_mp                      src/dafs.cpp:155-167
 *   calculate_similarity_score        src/dafs.cpp:713-764
 *   relax_basepairing_probability     src/dafs.cpp:326-375
 *   relax_matching_probability        src/dafs.cpp:258-324
 * Pinned bit-for-bit against the reference's own functions compiled from
 * /root/reference (oracle/_ref: ref_pct_run, tests/test_oracle_pct.py).
 *
 * Data layout = dafs_csr_batch: mp holds all i<j pairs row-major (rows = len[i]),
 * bp one unit per sequence; row_ptr blocks have rows+1 entries relative to nz_off. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CUTOFF 0.01 /* src/dafs.cpp:47 */

typedef struct { int rows; int* row_ptr; uint32_t* col; float* val; } csr_t;

static void csr_free(csr_t* m) { free(m->row_ptr); free(m->col); free(m->val); memset(m, 0, sizeof *m); }

static void csr_from(csr_t* m, int rows, const int* row_ptr, const uint32_t* col, const float* val) {
  const int n = row_ptr[rows];
  m->rows = rows;
  m->row_ptr = (int*)malloc(sizeof(int) * (rows + 1));
  m->col = (uint32_t*)malloc(sizeof(uint32_t) * (n > 0 ? n : 1));
  m->val = (float*)malloc(sizeof(float) * (n > 0 ? n : 1));
  memcpy(m->row_ptr, row_ptr, sizeof(int) * (rows + 1));
  memcpy(m->col, col, sizeof(uint32_t) * n);
  memcpy(m->val, val, sizeof(float) * n);
}

/* mp[i][i]: one entry (x, 1.0f) per row (src/align.cpp:41-44) */
static void csr_identity(csr_t* m, int rows) {
  m->rows = rows;
  m->row_ptr = (int*)malloc(sizeof(int) * (rows + 1));
  m->col = (uint32_t*)malloc(sizeof(uint32_t) * (rows > 0 ? rows : 1));
  m->val = (float*)malloc(sizeof(float) * (rows > 0 ? rows : 1));
  for (int r = 0; r < rows; ++r) { m->row_ptr[r] = r; m->col[r] = (uint32_t)r; m->val[r] = 1.0f; }
  m->row_ptr[rows] = rows;
}

/* transpose_mp, src/dafs.cpp:155-167: entries of column j collected in row order and
 * sorted by (row, value) -- rows are unique, so the order is ascending row. */
static void csr_transpose(const csr_t* m, int cols, csr_t* t) {
  const int n = m->row_ptr[m->rows];
  t->rows = cols;
  t->row_ptr = (int*)calloc((size_t)cols + 1, sizeof(int));
  t->col = (uint32_t*)malloc(sizeof(uint32_t) * (n > 0 ? n : 1));
  t->val = (float*)malloc(sizeof(float) * (n > 0 ? n : 1));
  for (int k = 0; k < n; ++k) ++t->row_ptr[m->col[k] + 1];
  for (int j = 0; j < cols; ++j) t->row_ptr[j + 1] += t->row_ptr[j];
  int* cur = (int*)malloc(sizeof(int) * (cols > 0 ? cols : 1));
  memcpy(cur, t->row_ptr, sizeof(int) * cols);
  for (int i = 0; i < m->rows; ++i)
    for (int k = m->row_ptr[i]; k < m->row_ptr[i + 1]; ++k) {
      const int p = cur[m->col[k]]++;
      t->col[p] = (uint32_t)i; t->val[p] = m->val[k];
    }
  free(cur);
}

/* calculate_similarity_score, src/dafs.cpp:713-764 */
static float similarity(const csr_t* mp, int L1, int L2) {
  float* dp = (float*)calloc((size_t)(L1 + 1) * (L2 + 1), sizeof(float));
  int* tr = (int*)calloc((size_t)(L1 + 1) * (L2 + 1), sizeof(int));
#define DP(i, j) dp[(size_t)(i) * (L2 + 1) + (j)]
#define TR(i, j) tr[(size_t)(i) * (L2 + 1) + (j)]
  for (int i = 1; i != L1 + 1; ++i) {
    int j = 1;
    for (int e = mp->row_ptr[i - 1]; e < mp->row_ptr[i]; ++e) {
      for (; j - 1 < (int)mp->col[e]; ++j) {
        DP(i, j) = DP(i, j - 1); TR(i, j) = TR(i, j - 1) + 1;
        if (DP(i, j) < DP(i - 1, j)) { DP(i, j) = DP(i - 1, j); TR(i, j) = TR(i - 1, j) + 1; }
      }
      DP(i, j) = DP(i - 1, j - 1) + mp->val[e]; TR(i, j) = TR(i - 1, j - 1) + 1;
      if (DP(i, j) < DP(i, j - 1)) { DP(i, j) = DP(i, j - 1); TR(i, j) = TR(i, j - 1) + 1; }
      if (DP(i, j) < DP(i - 1, j)) { DP(i, j) = DP(i - 1, j); TR(i, j) = TR(i - 1, j) + 1; }
      ++j;
    }
    for (; j < L2 + 1; ++j) {
      DP(i, j) = DP(i, j - 1); TR(i, j) = TR(i, j - 1) + 1;
      if (DP(i, j) < DP(i - 1, j)) { DP(i, j) = DP(i - 1, j); TR(i, j) = TR(i - 1, j) + 1; }
    }
  }
  const float s = DP(L1, L2) / TR(L1, L2);
#undef DP
#undef TR
  free(dp); free(tr);
  return s;
}

/* dense accumulator -> rows of entries with v = acc / sum_w > CUTOFF (dafs.cpp:302-311, 363-370) */
static void csr_from_dense(csr_t* m, const float* acc, int rows, int cols, float sum_w, int upper) {
  m->rows = rows;
  m->row_ptr = (int*)malloc(sizeof(int) * (rows + 1));
  int n = 0;
  for (int i = 0; i < rows; ++i)
    for (int j = upper ? i + 1 : 0; j < cols; ++j) if (acc[(size_t)i * cols + j] / sum_w > CUTOFF) ++n;
  m->col = (uint32_t*)malloc(sizeof(uint32_t) * (n > 0 ? n : 1));
  m->val = (float*)malloc(sizeof(float) * (n > 0 ? n : 1));
  n = 0;
  for (int i = 0; i < rows; ++i) {
    m->row_ptr[i] = n;
    for (int j = upper ? i + 1 : 0; j < cols; ++j) {
      const float v = acc[(size_t)i * cols + j] / sum_w;
      if (v > CUTOFF) { m->col[n] = (uint32_t)j; m->val[n] = v; ++n; }
    }
  }
  m->row_ptr[rows] = n;
}

static int g_n = 0;
static int* g_len = NULL;
static csr_t* g_mp = NULL; /* [n*n] */
static csr_t* g_bp = NULL; /* [n] */

static void pct_clear(void) {
  for (int k = 0; k < g_n * g_n; ++k) csr_free(&g_mp[k]);
  for (int k = 0; k < g_n; ++k) csr_free(&g_bp[k]);
  free(g_mp); free(g_bp); free(g_len);
  g_mp = NULL; g_bp = NULL; g_len = NULL; g_n = 0;
}

void oracle_pct_run(int n, const int* len, const long long* mp_row_off, const int* mp_row_ptr,
                    const long long* mp_nz_off, const uint32_t* mp_col, const float* mp_val,
                    const long long* bp_row_off, const int* bp_row_ptr, const long long* bp_nz_off,
                    const uint32_t* bp_col, const float* bp_val, float w_pct_a, float w_pct_s, float* sim_out,
                    long long* sizes) {
  pct_clear();
  const unsigned N = (unsigned)n;
  g_n = n;
  g_len = (int*)malloc(sizeof(int) * n);
  memcpy(g_len, len, sizeof(int) * n);
  g_mp = (csr_t*)calloc((size_t)n * n, sizeof(csr_t));
  g_bp = (csr_t*)calloc((size_t)n, sizeof(csr_t));
#define MP(a, b) g_mp[(size_t)(a) * n + (b)]
  long long u = 0;
  for (int i = 0; i < n; ++i) {
    csr_from(&g_bp[i], len[i], bp_row_ptr + bp_row_off[i], bp_col + bp_nz_off[i], bp_val + bp_nz_off[i]);
    csr_identity(&MP(i, i), len[i]);
    for (int j = i + 1; j < n; ++j, ++u) {
      csr_from(&MP(i, j), len[i], mp_row_ptr + mp_row_off[u], mp_col + mp_nz_off[u], mp_val + mp_nz_off[u]);
      csr_transpose(&MP(i, j), len[j], &MP(j, i)); /* dafs.cpp:1797-1799 */
    }
  }
  float* sim = (float*)malloc(sizeof(float) * (size_t)n * n); /* dafs.cpp:1812-1819 */
  for (int i = 0; i < n; ++i) {
    sim[(size_t)i * n + i] = 1.0;
    for (int j = i + 1; j < n; ++j) sim[(size_t)i * n + j] = sim[(size_t)j * n + i] = similarity(&MP(i, j), len[i], len[j]);
  }
  if (sim_out) memcpy(sim_out, sim, sizeof(float) * (size_t)n * n);

  if (w_pct_s != 0.0) { /* relax_basepairing_probability, dafs.cpp:326-375 */
    csr_t* nbp = (csr_t*)calloc((size_t)n, sizeof(csr_t));
    for (int x = 0; x < n; ++x) {
      const int L1 = len[x];
      float* p = (float*)calloc((size_t)L1 * L1, sizeof(float));
      float sum_w = 0.0;
      for (int y = 0; y < n; ++y) {
        float w = sim[(size_t)y * n + x];
        if (w_pct_s < 0.0) w *= 1.0 / N;
        else if (y == x) w *= 1.0 - w_pct_s;
        else w *= w_pct_s / (N - 1);
        sum_w += w;
        const csr_t* b = &g_bp[y];
        const csr_t* m = &MP(y, x);
        for (int k = 0; k < len[y]; ++k)
          for (int a = b->row_ptr[k]; a < b->row_ptr[k + 1]; ++a) {
            const int l = (int)b->col[a];
            const float p_kl = b->val[a];
            for (int c = m->row_ptr[k]; c < m->row_ptr[k + 1]; ++c) {
              const unsigned i = m->col[c];
              const float p_ik = m->val[c];
              for (int d = m->row_ptr[l]; d < m->row_ptr[l + 1]; ++d) {
                const unsigned j = m->col[d];
                const float p_jl = m->val[d];
                if (i < j) p[(size_t)i * L1 + j] += p_kl * p_ik * p_jl * w;
              }
            }
          }
      }
      csr_from_dense(&nbp[x], p, L1, L1, sum_w, 1);
      free(p);
    }
    for (int x = 0; x < n; ++x) csr_free(&g_bp[x]);
    free(g_bp);
    g_bp = nbp;
  }

  if (w_pct_a != 0.0) { /* relax_matching_probability, dafs.cpp:258-324 */
    csr_t* nmp = (csr_t*)calloc((size_t)n * n, sizeof(csr_t));
#define NMP(a, b) nmp[(size_t)(a) * n + (b)]
    for (int x = 0; x + 1 < n; ++x) {
      const int L1 = len[x];
      for (int y = x + 1; y < n; ++y) {
        const int L2 = len[y];
        float* post = (float*)calloc((size_t)L1 * L2, sizeof(float));
        float sum_w = 0.0;
        for (int z = 0; z < n; ++z) {
          float w = sim[(size_t)z * n + x] * sim[(size_t)z * n + y];
          if (w_pct_a < 0.0) w *= 1.0 / N;
          else if (z == x || z == y) w *= (1.0 - w_pct_a) / 2;
          else w *= w_pct_a / (N - 2);
          sum_w += w;
          const csr_t* a = &MP(z, x);
          const csr_t* b = &MP(z, y);
          for (int k = 0; k < len[z]; ++k)
            for (int c = a->row_ptr[k]; c < a->row_ptr[k + 1]; ++c) {
              const float p_ik = a->val[c];
              for (int d = b->row_ptr[k]; d < b->row_ptr[k + 1]; ++d)
                post[(size_t)a->col[c] * L2 + b->col[d]] += p_ik * b->val[d] * w;
            }
        }
        csr_from_dense(&NMP(x, y), post, L1, L2, sum_w, 0);
        csr_transpose(&NMP(x, y), L2, &NMP(y, x));
        free(post);
      }
    }
    for (int x = 0; x < n; ++x) csr_identity(&NMP(x, x), len[x]);
    for (int k = 0; k < n * n; ++k) csr_free(&g_mp[k]);
    free(g_mp);
    g_mp = nmp;
#undef NMP
  }
  free(sim);
  long long mr = 0, mn = 0, br = 0, bn = 0;
  for (int i = 0; i < n; ++i) {
    br += len[i] + 1; bn += g_bp[i].row_ptr[g_bp[i].rows];
    for (int j = i + 1; j < n; ++j) { mr += len[i] + 1; mn += MP(i, j).row_ptr[MP(i, j).rows]; }
  }
  sizes[0] = mr; sizes[1] = mn; sizes[2] = br; sizes[3] = bn;
}

void oracle_pct_copy(int* mp_row_ptr, uint32_t* mp_col, float* mp_val, int* bp_row_ptr, uint32_t* bp_col, float* bp_val) {
  const int n = g_n;
  long long rp = 0, nz = 0;
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j) {
      const csr_t* m = &MP(i, j);
      const int k = m->row_ptr[m->rows];
      memcpy(mp_row_ptr + rp, m->row_ptr, sizeof(int) * (m->rows + 1));
      memcpy(mp_col + nz, m->col, sizeof(uint32_t) * k);
      memcpy(mp_val + nz, m->val, sizeof(float) * k);
      rp += m->rows + 1; nz += k;
    }
  rp = 0; nz = 0;
  for (int i = 0; i < n; ++i) {
    const csr_t* m = &g_bp[i];
    const int k = m->row_ptr[m->rows];
    memcpy(bp_row_ptr + rp, m->row_ptr, sizeof(int) * (m->rows + 1));
    memcpy(bp_col + nz, m->col, sizeof(uint32_t) * k);
    memcpy(bp_val + nz, m->val, sizeof(float) * k);
    rp += m->rows + 1; nz += k;
  }
}
#undef MP
