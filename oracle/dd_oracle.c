/* TEST INFRASTRUCTURE ONLY -- the product (libdafs_cuda) never links, loads or
 * calls anything in oracle/.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it, as the checker.
 *
 * Plain-C restatement of the dual-decomposition path of satoken/dafs:
 *   SparseNussinov::decode(w,p,q,ss)          src/nussinov.cpp:207-298
 *   SparseNussinov::decode(p,ss,str)          src/nussinov.cpp:300-392  (w=1,q=0)
 *   SparseNeedlemanWunsch::initialize         src/needleman_wunsch.cpp:198-253
 *   SparseNeedlemanWunsch::decode(p,q,al)     src/needleman_wunsch.cpp:255-338
 *   SparseNeedlemanWunsch::decode(p,al)       src/needleman_wunsch.cpp:340-422
 *   DAFS::solve_by_dd                         src/dafs.cpp:1006-1295
 *     (SPARSE_UPDATE on, ADAGRAD/ADAM off: src/dafs.cpp:67-69)
 * Parity: PINNED.  tests/test_oracle_dd.py checks every function here
 * bit-for-bit against oracle/_ref (the reference's own translation units,
 * compiled from /root/reference) and against the fixtures in tests/golden/
 * that were generated from oracle/_ref.
 *
 * Scalar FP32, every operation rounded on its own (-ffp-contract=off), like
 * the reference's x86-64 SSE2 build; matrices are dense row-major like VVF.
 */
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define NOPAIR 0xffffffffu   /* -1u */

/* ---- SparseNussinov::decode (nussinov.cpp:207-298) ------------------------
 * p, q: L x L row-major, only i<j read.  q == NULL: the plain overload
 * (:300-392), s = p - th.  Returns dp[0][L-1]. */
float oracle_nussinov(int L, float th, float w, const float* p, const float* q, uint32_t* ss) {
  float* dp = (float*)calloc((size_t)L * L, sizeof(float));
  uint32_t* tr = (uint32_t*)calloc((size_t)L * L, sizeof(uint32_t));
  /* bp[j]: candidate list of column j in push order (descending i) */
  int* cnt = (int*)calloc(L, sizeof(int));
  int* ck = (int*)malloc(sizeof(int) * (size_t)L * L);
  float* cs = (float*)malloc(sizeof(float) * (size_t)L * L);
  for (int l = 1; l < L; ++l) {
    for (int i = 0; i + l < L; ++i) {
      const int j = i + l;
      float v = 0.0f;
      int t = 0;
      if (i + 1 < j) { v = dp[(size_t)(i + 1) * L + j]; t = 1; }
      if (i < j - 1 && v < dp[(size_t)i * L + j - 1]) { v = dp[(size_t)i * L + j - 1]; t = 2; }
      if (i + 1 < j - 1) {
        float s;
        if (q) s = w * (p[(size_t)i * L + j] - th) - q[(size_t)i * L + j];
        else s = p[(size_t)i * L + j] - th;
        if (s > 0.0) {
          const float c = dp[(size_t)(i + 1) * L + j - 1] + s;
          ck[(size_t)j * L + cnt[j]] = i;
          cs[(size_t)j * L + cnt[j]] = c;
          cnt[j]++;
          if (v < c) { v = c; t = 3; }
        }
      }
      for (int e = 0; e < cnt[j]; ++e) {
        const int k = ck[(size_t)j * L + e];
        const float s = cs[(size_t)j * L + e];
        if (i < k) {
          const float c = dp[(size_t)i * L + k - 1] + s;
          if (v < c) { v = c; t = k - i + 3; }
        }
      }
      dp[(size_t)i * L + j] = v;
      tr[(size_t)i * L + j] = (uint32_t)t;
    }
  }
  /* traceback, :265-295 (explicit stack; order does not affect ss) */
  for (int i = 0; i < L; ++i) ss[i] = NOPAIR;
  int* st = (int*)malloc(sizeof(int) * 2 * (size_t)(2 * L + 4));
  int sp = 0;
  st[0] = 0; st[1] = L - 1; sp = 1;
  while (sp > 0) {
    --sp;
    const int i = st[2 * sp], j = st[2 * sp + 1];
    const uint32_t t = tr[(size_t)i * L + j];
    switch (t) {
      case 0: break;
      case 1: st[2 * sp] = i + 1; st[2 * sp + 1] = j; ++sp; break;
      case 2: st[2 * sp] = i; st[2 * sp + 1] = j - 1; ++sp; break;
      case 3: ss[i] = (uint32_t)j; st[2 * sp] = i + 1; st[2 * sp + 1] = j - 1; ++sp; break;
      default: {
        const int k = i + (int)t - 3;
        st[2 * sp] = i; st[2 * sp + 1] = k - 1; ++sp;
        ss[k] = (uint32_t)j;
        st[2 * sp] = k + 1; st[2 * sp + 1] = j - 1; ++sp;
        break;
      }
    }
  }
  const float ret = dp[L - 1];
  free(st); free(cs); free(ck); free(cnt); free(tr); free(dp);
  return ret;
}

/* ---- SparseNeedlemanWunsch::initialize (needleman_wunsch.cpp:198-253) -----
 * env: 2*(L1+1) uints {first, second}. */
void oracle_nw_envelope(int L1, int L2, float th, const float* p, uint32_t* env) {
  uint32_t* first = env;          /* interleaved below */
  (void)first;
  for (int i = 0; i <= L1; ++i) { env[2 * i] = 0; env[2 * i + 1] = 0; }
  for (int i = 1; i <= L1; ++i) {
    for (int k = 1; k != L2 + 1; ++k) {
      if (p[(size_t)(i - 1) * L2 + k - 1] - th >= 0.0) {
        if ((uint32_t)(k - 1) < env[2 * (i - 1)]) env[2 * (i - 1)] = (uint32_t)(k - 1);
        env[2 * i] = (uint32_t)k;
        break;
      }
    }
    if (env[2 * i] == 0) {
      env[2 * i] = env[2 * (i - 1)];
      env[2 * i + 1] = env[2 * (i - 1) + 1];
      continue;
    }
    for (int k = L2; k != 0; --k) {
      if (p[(size_t)(i - 1) * L2 + k - 1] - th >= 0.0) {
        if ((uint32_t)(k - 1) > env[2 * (i - 1) + 1]) env[2 * (i - 1) + 1] = (uint32_t)(k - 1);
        env[2 * i + 1] = (uint32_t)k;
        break;
      }
    }
  }
  env[2 * L1 + 1] = (uint32_t)L2;
  {
    uint32_t v = (uint32_t)L2;
    for (int i = L1; i != 0; --i) { if (env[2 * i] < v) v = env[2 * i]; env[2 * i] = v; }
  }
  {
    uint32_t v = 0;
    for (int i = 0; i != L1 + 1; ++i) { if (env[2 * i + 1] > v) v = env[2 * i + 1]; env[2 * i + 1] = v; }
  }
  for (int i = 1; i != L1 + 1; ++i)
    if (env[2 * (i - 1) + 1] < env[2 * i]) env[2 * i] = env[2 * (i - 1) + 1];
}

/* ---- SparseNeedlemanWunsch::decode (needleman_wunsch.cpp:255-338, :340-422) -
 * q == NULL: the plain overload.  Returns dp[L1][L2]; *bad is set when the
 * traceback meets an unfilled cell (the reference would spin there: its
 * default arm is an assert that -DNDEBUG removes, :309). */
float oracle_nw_decode(int L1, int L2, float th, const float* p, const float* q,
                       const uint32_t* env, uint32_t* al, int* bad) {
  const size_t W = (size_t)L2 + 1;
  float* dp = (float*)malloc(sizeof(float) * (L1 + 1) * W);
  char* tr = (char*)malloc((L1 + 1) * W);
  for (size_t c = 0; c < (L1 + 1) * W; ++c) { dp[c] = -FLT_MAX; tr[c] = ' '; }
  dp[0] = 0.0f;
  for (int i = 1; i <= L1; ++i) { dp[i * W] = 0.0f; tr[i * W] = 'X'; }
  for (int k = 1; k <= L2; ++k) { dp[k] = 0.0f; tr[k] = 'Y'; }
  for (int i = 1; i <= L1; ++i) {
    for (uint32_t k = env[2 * i]; k <= env[2 * i + 1]; ++k) {
      if (k == 0) continue;
      float v;
      if (q) v = dp[(i - 1) * W + k - 1] + p[(size_t)(i - 1) * L2 + k - 1] - th + q[(size_t)(i - 1) * L2 + k - 1];
      else v = dp[(i - 1) * W + k - 1] + p[(size_t)(i - 1) * L2 + k - 1] - th;
      char t = 'M';
      if (v < dp[(i - 1) * W + k]) { v = dp[(i - 1) * W + k]; t = 'X'; }
      if (v < dp[i * W + k - 1]) { v = dp[i * W + k - 1]; t = 'Y'; }
      dp[i * W + k] = v;
      tr[i * W + k] = t;
    }
  }
  if (bad) *bad = 0;
  for (int i = 0; i < L1; ++i) al[i] = NOPAIR;
  {
    int i = L1, k = L2;
    while (i > 0 || k > 0) {
      const char t = tr[i * W + k];
      if (t == 'M') { --i; --k; al[i] = (uint32_t)k; }
      else if (t == 'X') { --i; }
      else if (t == 'Y') { --k; }
      else { if (bad) *bad = 1; break; }
    }
  }
  const float ret = dp[L1 * W + L2];
  free(tr); free(dp);
  return ret;
}

/* ---- DAFS::solve_by_dd (dafs.cpp:1006-1295) --------------------------------
 * p_x: L1 x L1, p_y: L2 x L2, p_z: L1 x L2 dense row-major.  th_s is th_s_[0]
 * (the threshold of s_decoder_, dafs.cpp:1759), min_th_s = min(th_s_) (:1018).
 * Outputs x[L1], y[L2], z[L1]; *iters = t at exit, *violated, *ncbp.
 * trace (may be NULL): per iteration {s, eta_used, violated} as 3 floats,
 * capacity trace_cap iterations (what -v 2 prints, :1273).               */
#define CUTOFF 0.01

typedef struct { uint32_t i, j, k, l; } cbp_t;

static int cmp_u32(const void* a, const void* b) {
  const uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
  return x < y ? -1 : x > y;
}

/* sorted-unique per-row projection lists (:1046-1061) in CSR form */
static void build_proj(int rows, const uint32_t* r, const uint32_t* c, size_t n, int** ptr_out,
                       uint32_t** col_out) {
  int* ptr = (int*)calloc(rows + 1, sizeof(int));
  for (size_t u = 0; u < n; ++u) ptr[r[u] + 1]++;
  for (int i = 0; i < rows; ++i) ptr[i + 1] += ptr[i];
  uint32_t* col = (uint32_t*)malloc(sizeof(uint32_t) * (n ? n : 1));
  int* fill = (int*)malloc(sizeof(int) * (rows + 1));
  memcpy(fill, ptr, sizeof(int) * (rows + 1));
  for (size_t u = 0; u < n; ++u) col[fill[r[u]]++] = c[u];
  int* uptr = (int*)calloc(rows + 1, sizeof(int));
  int w = 0;
  for (int i = 0; i < rows; ++i) {
    qsort(col + ptr[i], ptr[i + 1] - ptr[i], sizeof(uint32_t), cmp_u32);
    uptr[i] = w;
    for (int e = ptr[i]; e < ptr[i + 1]; ++e)
      if (e == ptr[i] || col[e] != col[e - 1]) col[w++] = col[e];
  }
  uptr[rows] = w;
  free(fill); free(ptr);
  *ptr_out = uptr; *col_out = col;
}

float oracle_dd_solve(int L1, int L2, int N1, int N2, const float* p_x, const float* p_y,
                      const float* p_z, float w_, float th_s, float min_th_s, float th_a,
                      float eta0, int t_max, uint32_t* x, uint32_t* y, uint32_t* z, int* iters,
                      uint32_t* violated_out, int64_t* ncbp_out, float* trace, int trace_cap) {
  const unsigned uN1 = (unsigned)N1, uN2 = (unsigned)N2;
  /* CBP enumeration, :1016-1045 (float > double compares against CUTOFF) */
  size_t cap = 1024, n = 0;
  cbp_t* cbp = (cbp_t*)malloc(sizeof(cbp_t) * cap);
  for (int i = 0; i + 1 < L1; ++i)
    for (int j = i + 1; j < L1; ++j)
      if (p_x[(size_t)i * L1 + j] > CUTOFF)
        for (int k = 0; k + 1 < L2; ++k)
          if (p_z[(size_t)i * L2 + k] > CUTOFF)
            for (int l = k + 1; l < L2; ++l)
              if (p_y[(size_t)k * L2 + l] > CUTOFF && p_z[(size_t)j * L2 + l] > CUTOFF) {
                const float p = (uN1 * p_x[(size_t)i * L1 + j] + uN2 * p_y[(size_t)k * L2 + l]) / (uN1 + uN2);
                const float q = (p_z[(size_t)i * L2 + k] + p_z[(size_t)j * L2 + l]) / 2;
                if (p - min_th_s > 0.0 && w_ * (p - min_th_s) + (q - th_a) > 0.0) {
                  if (n == cap) { cap *= 2; cbp = (cbp_t*)realloc(cbp, sizeof(cbp_t) * cap); }
                  cbp[n].i = i; cbp[n].j = j; cbp[n].k = k; cbp[n].l = l; ++n;
                }
              }
  uint32_t *ra = (uint32_t*)malloc(sizeof(uint32_t) * (2 * n + 1)), *ca = (uint32_t*)malloc(sizeof(uint32_t) * (2 * n + 1));
  int *cx_ptr, *cy_ptr, *cz_ptr; uint32_t *cx, *cy, *cz;
  for (size_t u = 0; u < n; ++u) { ra[u] = cbp[u].i; ca[u] = cbp[u].j; }
  build_proj(L1, ra, ca, n, &cx_ptr, &cx);
  for (size_t u = 0; u < n; ++u) { ra[u] = cbp[u].k; ca[u] = cbp[u].l; }
  build_proj(L2, ra, ca, n, &cy_ptr, &cy);
  for (size_t u = 0; u < n; ++u) { ra[2 * u] = cbp[u].i; ca[2 * u] = cbp[u].k; ra[2 * u + 1] = cbp[u].j; ca[2 * u + 1] = cbp[u].l; }
  build_proj(L1, ra, ca, 2 * n, &cz_ptr, &cz);
  free(ra); free(ca);

  uint32_t* env = (uint32_t*)malloc(sizeof(uint32_t) * 2 * (L1 + 1));
  oracle_nw_envelope(L1, L2, th_a, p_z, env);                         /* :1064 */

  float* q_x = (float*)calloc((size_t)L1 * L1, sizeof(float));
  float* q_y = (float*)calloc((size_t)L2 * L2, sizeof(float));
  float* q_z = (float*)calloc((size_t)L1 * L2, sizeof(float));
  int* t_x = (int*)malloc(sizeof(int) * (size_t)L1 * L1);
  int* t_y = (int*)malloc(sizeof(int) * (size_t)L2 * L2);
  int* t_z = (int*)malloc(sizeof(int) * (size_t)L1 * L2);
  const float w_x = w_ * 2 * uN1 / (uN1 + uN2), w_y = w_ * 2 * uN2 / (uN1 + uN2);   /* :1091-1092 */
  float c = 0.0f;
  float eta = eta0;
  float s_prev = 0.0f;
  unsigned violated = 0;
  int t;
  for (t = 0; t != t_max; ++t) {
    float s = 0.0f;
    s += oracle_nussinov(L1, th_s, w_x, p_x, q_x, x);
    s += oracle_nussinov(L2, th_s, w_y, p_y, q_y, y);
    int bad = 0;
    s += oracle_nw_decode(L1, L2, th_a, p_z, q_z, env, z, &bad);
    violated = 0;
    memset(t_x, 0, sizeof(int) * (size_t)L1 * L1);
    memset(t_y, 0, sizeof(int) * (size_t)L2 * L2);
    memset(t_z, 0, sizeof(int) * (size_t)L1 * L2);
    for (size_t u = 0; u < n; ++u) {                                   /* :1103-1117 */
      const uint32_t i = cbp[u].i, j = cbp[u].j, k = cbp[u].k, l = cbp[u].l;
      const float s_w = q_x[(size_t)i * L1 + j] + q_y[(size_t)k * L2 + l] - q_z[(size_t)i * L2 + k] - q_z[(size_t)j * L2 + l];
      if (s_w > 0.0f) {
        s += s_w;
        t_x[(size_t)i * L1 + j]++; t_y[(size_t)k * L2 + l]++;
        t_z[(size_t)i * L2 + k]++; t_z[(size_t)j * L2 + l]++;
      }
    }
    for (int i = 0; i < L1; ++i) {                                     /* :1121-1150 */
      const uint32_t j = x[i];
      if (j != NOPAIR && t_x[(size_t)i * L1 + j] != 1) {
        violated++;
        q_x[(size_t)i * L1 + j] -= eta * (t_x[(size_t)i * L1 + j] - 1);
      }
      for (int e = cx_ptr[i]; e < cx_ptr[i + 1]; ++e) {
        const uint32_t jj = cx[e];
        if (x[i] != jj && t_x[(size_t)i * L1 + jj] != 0) {
          violated++;
          q_x[(size_t)i * L1 + jj] -= eta * t_x[(size_t)i * L1 + jj];
        }
      }
    }
    for (int k = 0; k < L2; ++k) {                                     /* :1172-1201 */
      const uint32_t l = y[k];
      if (l != NOPAIR && t_y[(size_t)k * L2 + l] != 1) {
        violated++;
        q_y[(size_t)k * L2 + l] -= eta * (t_y[(size_t)k * L2 + l] - 1);
      }
      for (int e = cy_ptr[k]; e < cy_ptr[k + 1]; ++e) {
        const uint32_t ll = cy[e];
        if (y[k] != ll && t_y[(size_t)k * L2 + ll] != 0) {
          violated++;
          q_y[(size_t)k * L2 + ll] -= eta * t_y[(size_t)k * L2 + ll];
        }
      }
    }
    for (int i = 0; i < L1; ++i) {                                     /* :1223-1254 */
      const uint32_t k = z[i];
      if (k != NOPAIR) {
        if (t_z[(size_t)i * L2 + k] > 1) violated++;
        const float nv = q_z[(size_t)i * L2 + k] - eta * (1 - t_z[(size_t)i * L2 + k]);
        q_z[(size_t)i * L2 + k] = (0.0f < nv) ? nv : 0.0f;             /* std::max(0.0f, nv) */
      }
      for (int e = cz_ptr[i]; e < cz_ptr[i + 1]; ++e) {
        const uint32_t kk = cz[e];
        if (z[i] != kk) {
          if (t_z[(size_t)i * L2 + kk] > 0) violated++;
          const float nv = q_z[(size_t)i * L2 + kk] + eta * t_z[(size_t)i * L2 + kk];
          q_z[(size_t)i * L2 + kk] = (0.0f < nv) ? nv : 0.0f;
        }
      }
    }
    if (trace && t < trace_cap) { trace[3 * t] = s; trace[3 * t + 1] = eta; trace[3 * t + 2] = (float)violated; }
    if (violated == 0) break;                                          /* :1278 */
    if (s > s_prev || t == 0) {                                        /* :1283-1288 */
      const float a = 4.0f * n - violated;
      c += ((0.0f < a) ? a : 0.0f) / (4.0 * n);
      eta = eta0 / (1.0 + c);
    }
    s_prev = s;
  }
  if (iters) *iters = t;
  if (violated_out) *violated_out = violated;
  if (ncbp_out) *ncbp_out = (int64_t)n;
  free(t_z); free(t_y); free(t_x); free(q_z); free(q_y); free(q_x); free(env);
  free(cz); free(cz_ptr); free(cy); free(cy_ptr); free(cx); free(cx_ptr); free(cbp);
  return s_prev;
}
