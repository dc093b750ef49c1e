/* TEST INFRASTRUCTURE ONLY -- the product (libdafs_cuda) never links, loads or
 * calls anything in oracle/.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it, as the checker.
 *
 * Plain-C restatement of the ProbCons pair-HMM posterior path of satoken/dafs
 * (NumInsertStates = 1: states 0 = M, 1 = X consumes seq1, 2 = Y consumes seq2).
 * Parity: PINNED.  tests/test_oracle_pairhmm.py checks this file bit-for-bit
 * against oracle/_ref (the reference's own ProbCons::calculate compiled from
 * /root/reference) and against the README guide-tree scores (README.md:59).
 *
 * All arithmetic is scalar FP32 with every multiply and add rounded
 * separately (build with -ffp-contract=off), which is what the reference's
 * x86-64 build does; EXP is evaluated in double like the reference.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define LOG_ZERO (-2e20f)          /* probconsRNA/ScoreType.h:18  */
#define LOG_UNDERFLOW 7.5f         /* probconsRNA/ScoreType.h:178 */

/* ---- parameters: probconsRNA/Defaults.h:19-39 -------------------------- */
static const float k_init[3] = {0.9588437676f, 0.0205782652f, 0.0205782652f};
static const float k_gap_open = 0.0190259293f, k_gap_extend = 0.3269913495f;
static const float k_emit_single[6] = {0.2270790040f, 0.2422080040f, 0.2839320004f,
                                       0.2464679927f, 0.2464679927f, 0.0003124650f};
static const float k_emit_pairs[6][6] = {
    {0.1487240046f, 0.0184142999f, 0.0361397006f, 0.0238473993f, 0.0238473993f, 0.0000375308f},
    {0.0184142999f, 0.1583919972f, 0.0275536999f, 0.0389291011f, 0.0389291011f, 0.0000815823f},
    {0.0361397006f, 0.0275536999f, 0.1979320049f, 0.0244289003f, 0.0244289003f, 0.0000824765f},
    {0.0238473993f, 0.0389291011f, 0.0244289003f, 0.1557479948f, 0.1557479948f, 0.0000743985f},
    {0.0238473993f, 0.0389291011f, 0.0244289003f, 0.1557479948f, 0.1557479948f, 0.0000743985f},
    {0.0000375308f, 0.0000815823f, 0.0000824765f, 0.0000743985f, 0.0000743985f, 0.0000263252f}};

/* Symbol classes.  The reference indexes 256x256 tables by raw chars
 * (ProbabilisticModel.h:43-44) filled case-insensitively over "ACGUTN"
 * (wrapper.cpp:134-171); everything else keeps 1e-10 / 1e-5.  Six classes
 * 0..5 = A C G U T N and class 6 = "other" carry the same information. */
int oracle_symbol_class(unsigned char c) {
  switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'U': case 'u': return 3;
    case 'T': case 't': return 4;
    case 'N': case 'n': return 5;
    default: return 6;
  }
}

typedef struct {
  float init[3];      /* log initialDistribution            (ProbabilisticModel.h:76) */
  float trans[3][3];  /* log transProb                      (:77-78)                  */
  float ins[7];       /* log insProb per symbol class       (:83-84)                  */
  float match[7][7];  /* log matchProb per class pair       (:85-86)                  */
} pairhmm_tables;

/* LOG() is log(float) under "using namespace std" -> the float overload. */
static float LOGf(float x) { return logf(x); }

void oracle_pairhmm_tables(pairhmm_tables* t) {
  float tm[3][3] = {{0}};
  /* ProbabilisticModel.h:58-72 */
  tm[0][0] = 1;
  tm[0][1] = k_gap_open;
  tm[0][2] = k_gap_open;
  tm[0][0] -= (k_gap_open + k_gap_open);
  tm[1][1] = k_gap_extend;
  tm[2][2] = k_gap_extend;
  tm[1][0] = 1 - k_gap_extend;
  tm[2][0] = 1 - k_gap_extend;
  for (int i = 0; i < 3; i++) {
    t->init[i] = LOGf(k_init[i]);
    for (int j = 0; j < 3; j++) t->trans[i][j] = LOGf(tm[i][j]);
  }
  for (int a = 0; a < 7; a++) {
    t->ins[a] = LOGf(a < 6 ? k_emit_single[a] : (float)1e-5);
    for (int b = 0; b < 7; b++)
      t->match[a][b] = LOGf((a < 6 && b < 6) ? k_emit_pairs[a][b] : (float)1e-10);
  }
}

/* ---- ScoreType.h:187-198 ------------------------------------------------ */
static inline float LOOKUP(float x) {
  if (x <= 1.00f) return ((-0.009350833524763f * x + 0.130659527668286f) * x + 0.498799810682272f) * x + 0.693203116424741f;
  if (x <= 2.50f) return ((-0.014532321752540f * x + 0.139942324101744f) * x + 0.495635523139337f) * x + 0.692140569840976f;
  if (x <= 4.50f) return ((-0.004605031767994f * x + 0.063427417320019f) * x + 0.695956496475118f) * x + 0.514272634594009f;
  return ((-0.000458661602210f * x + 0.009695946122598f) * x + 0.930734667215156f) * x + 0.168037164329057f;
}

/* ScoreType.h:233-238 and :259-262 */
static inline float LOG_ADD(float x, float y) {
  if (x < y) return (x == LOG_ZERO || y - x >= LOG_UNDERFLOW) ? y : LOOKUP(y - x) + x;
  return (y == LOG_ZERO || x - y >= LOG_UNDERFLOW) ? x : LOOKUP(x - y) + y;
}
#define LOG_PLUS_EQUALS(x, y) ((x) = LOG_ADD((x), (y)))

/* ScoreType.h:37-57: double coefficients, so the polynomial runs in double
 * and is rounded to float on return. */
static inline float EXPf(float xf) {
  double x = xf;
  if (x > -2) {
    if (x > -0.5) {
      if (x > 0) return (float)exp(x);
      return (float)((((0.03254409303190190000 * x + 0.16280432765779600000) * x + 0.49929760485974900000) * x + 0.99995149601363700000) * x + 0.99999925508501600000);
    }
    if (x > -1)
      return (float)((((0.01973899026052090000 * x + 0.13822379685007000000) * x + 0.48056651562365000000) * x + 0.99326940370383500000) * x + 0.99906756856399500000);
    return (float)((((0.00940528203591384000 * x + 0.09414963667859410000) * x + 0.40825793595877300000) * x + 0.93933625499130400000) * x + 0.98369508190545300000);
  }
  if (x > -8) {
    if (x > -4)
      return (float)((((0.00217245711583303000 * x + 0.03484829428350620000) * x + 0.22118199801337800000) * x + 0.67049462206469500000) * x + 0.83556950223398500000);
    return (float)((((0.00012398771025456900 * x + 0.00349155785951272000) * x + 0.03727721426017900000) * x + 0.17974997741536900000) * x + 0.33249299994217400000);
  }
  if (x > -16)
    return (float)((((0.00000051741713416603 * x + 0.00002721456879608080) * x + 0.00053418601865636800) * x + 0.00464101989351936000) * x + 0.01507447981459420000);
  return 0;
}

/* ---- forward: ProbabilisticModel.h:105-179 ------------------------------
 * F is [(L1+1)*(L2+1)][3], state-minor, like the reference's VF. */
static void forward(const pairhmm_tables* T, const int* c1, int L1, const int* c2, int L2, float* F) {
  const int W = L2 + 1;
  for (long n = 0; n < 3L * (L1 + 1) * W; n++) F[n] = LOG_ZERO;
  if (L1 >= 1 && L2 >= 1) F[0 + 3 * (1 * W + 1)] = T->init[0] + T->match[c1[1]][c2[1]];
  if (L1 >= 1) F[1 + 3 * (1 * W + 0)] = T->init[1] + T->ins[c1[1]];
  if (L2 >= 1) F[2 + 3 * (0 * W + 1)] = T->init[2] + T->ins[c2[1]];
  for (int i = 0; i <= L1; i++)
    for (int j = 0; j <= L2; j++) {
      if (!(i > 1 || j > 1)) continue;
      float* f = F + 3 * ((long)i * W + j);
      if (i > 0 && j > 0) {
        const float* d = F + 3 * ((long)(i - 1) * W + (j - 1));
        f[0] = d[0] + T->trans[0][0];
        LOG_PLUS_EQUALS(f[0], d[1] + T->trans[1][0]);
        LOG_PLUS_EQUALS(f[0], d[2] + T->trans[2][0]);
        f[0] += T->match[c1[i]][c2[j]];
      }
      if (i > 0) {
        const float* u = F + 3 * ((long)(i - 1) * W + j);
        f[1] = T->ins[c1[i]] + LOG_ADD(u[0] + T->trans[0][1], u[1] + T->trans[1][1]);
      }
      if (j > 0) {
        const float* l = F + 3 * ((long)i * W + (j - 1));
        f[2] = T->ins[c2[j]] + LOG_ADD(l[0] + T->trans[0][2], l[2] + T->trans[2][2]);
      }
    }
}

/* ---- backward: ProbabilisticModel.h:197-259 ------------------------------ */
static void backward(const pairhmm_tables* T, const int* c1, int L1, const int* c2, int L2, float* B) {
  const int W = L2 + 1;
  for (long n = 0; n < 3L * (L1 + 1) * W; n++) B[n] = LOG_ZERO;
  for (int k = 0; k < 3; k++) B[3 * ((long)(L1 + 1) * W - 1) + k] = T->init[k];
  for (int i = L1; i >= 0; i--)
    for (int j = L2; j >= 0; j--) {
      float* b = B + 3 * ((long)i * W + j);
      if (i < L1 && j < L2) {
        const float pxy = B[0 + 3 * ((long)(i + 1) * W + (j + 1))] + T->match[c1[i + 1]][c2[j + 1]];
        for (int k = 0; k < 3; k++) LOG_PLUS_EQUALS(b[k], pxy + T->trans[k][0]);
      }
      if (i < L1) {
        const float bx = B[1 + 3 * ((long)(i + 1) * W + j)];
        LOG_PLUS_EQUALS(b[0], bx + T->ins[c1[i + 1]] + T->trans[0][1]);
        LOG_PLUS_EQUALS(b[1], bx + T->ins[c1[i + 1]] + T->trans[1][1]);
      }
      if (j < L2) {
        const float by = B[2 + 3 * ((long)i * W + (j + 1))];
        LOG_PLUS_EQUALS(b[0], by + T->ins[c2[j + 1]] + T->trans[0][2]);
        LOG_PLUS_EQUALS(b[2], by + T->ins[c2[j + 1]] + T->trans[2][2]);
      }
    }
}

/* ---- total: ProbabilisticModel.h:337-365 --------------------------------- */
static float total_probability(int L1, int L2, const float* F, const float* B) {
  const int W = L2 + 1;
  const long last = 3 * ((long)(L1 + 1) * W - 1);
  float tf = LOG_ZERO, tb;
  for (int k = 0; k < 3; k++) LOG_PLUS_EQUALS(tf, F[k + last] + B[k + last]);
  tb = F[0 + 3 * (1 * W + 1)] + B[0 + 3 * (1 * W + 1)];
  LOG_PLUS_EQUALS(tb, F[1 + 3 * (1 * W + 0)] + B[1 + 3 * (1 * W + 0)]);
  LOG_PLUS_EQUALS(tb, F[2 + 3 * (0 * W + 1)] + B[2 + 3 * (0 * W + 1)]);
  return (tf + tb) / 2;
}

/* Scratch-owning single-pair driver.
 * posterior: ProbabilisticModel.h:374-403; th zeroing: wrapper.cpp:125-128;
 * strict > th sparsification: align.cpp:69-78.
 * dense (may be NULL): L1*L2 floats, 0-based, zero where dropped.
 * dbg (may be NULL): [0]=total.  Returns nnz, or -1 on allocation failure. */
int oracle_pairhmm_pair(const char* s1, int L1, const char* s2, int L2, float th,
                        float* dense, int* row_ptr, unsigned* col, float* val, float* dbg) {
  static __thread pairhmm_tables T;
  static __thread int have = 0;
  if (!have) { oracle_pairhmm_tables(&T); have = 1; }
  if (L1 < 1 || L2 < 1) return -1; /* the reference indexes iter[1] unconditionally */
  const int W = L2 + 1;
  int* c1 = (int*)malloc(sizeof(int) * (L1 + 2));
  int* c2 = (int*)malloc(sizeof(int) * (L2 + 2));
  float* F = (float*)malloc(sizeof(float) * 3 * (size_t)(L1 + 1) * W);
  float* B = (float*)malloc(sizeof(float) * 3 * (size_t)(L1 + 1) * W);
  if (!c1 || !c2 || !F || !B) { free(c1); free(c2); free(F); free(B); return -1; }
  c1[0] = c2[0] = 6; /* '@' at index 0 (wrapper.cpp:105-113); never read for emission */
  for (int i = 0; i < L1; i++) c1[i + 1] = oracle_symbol_class((unsigned char)s1[i]);
  for (int j = 0; j < L2; j++) c2[j + 1] = oracle_symbol_class((unsigned char)s2[j]);
  forward(&T, c1, L1, c2, L2, F);
  backward(&T, c1, L1, c2, L2, B);
  const float total = total_probability(L1, L2, F, B);
  if (dbg) dbg[0] = total;
  int nnz = 0;
  if (row_ptr) row_ptr[0] = 0;
  for (int i = 1; i <= L1; i++) {
    for (int j = 1; j <= L2; j++) {
      const long ij = 3 * ((long)i * W + j);
      float x = F[ij] + B[ij] - total;
      float p = EXPf(x < 0.0f ? x : 0.0f);
      p = (p >= th) ? p : 0.0f;
      const int keep = p > th;
      if (dense) dense[(long)(i - 1) * L2 + (j - 1)] = keep ? p : 0.0f;
      if (keep) {
        if (col) col[nnz] = (unsigned)(j - 1);
        if (val) val[nnz] = p;
        nnz++;
      }
    }
    if (row_ptr) row_ptr[i] = nnz;
  }
  free(c1); free(c2); free(F); free(B);
  return nnz;
}

/* ---- all-pairs driver (align.cpp:35-52) with an optional pthread fan-out --
 * Used only for parity at small N and for the timed CPU baseline in bench.py.
 * nnz_out[p] for pair p in row-major i<j order.  first/count select a
 * sub-range of pairs (bounded sample). */
typedef struct {
  int nseq; const char* const* seq; const int* len; float th;
  long first, count; int* nnz_out; volatile long* next;
} allpairs_job;

static void pair_from_index(long p, int n, int* i, int* j) {
  int a = 0; long rem = p;
  while (rem >= n - 1 - a) { rem -= n - 1 - a; a++; }
  *i = a; *j = a + 1 + (int)rem;
}

static void* allpairs_worker(void* arg) {
  allpairs_job* J = (allpairs_job*)arg;
  for (;;) {
    long k = __sync_fetch_and_add(J->next, 1);
    if (k >= J->count) break;
    int i, j;
    pair_from_index(J->first + k, J->nseq, &i, &j);
    int nnz = oracle_pairhmm_pair(J->seq[i], J->len[i], J->seq[j], J->len[j], J->th,
                                  NULL, NULL, NULL, NULL, NULL);
    if (J->nnz_out) J->nnz_out[k] = nnz;
  }
  return NULL;
}

int oracle_pairhmm_all(int nseq, const char* const* seq, const int* len, float th,
                       long first, long count, int nthreads, int* nnz_out) {
  volatile long next = 0;
  allpairs_job J = {nseq, seq, len, th, first, count, nnz_out, &next};
  if (nthreads <= 1) { allpairs_worker(&J); return 0; }
  pthread_t* t = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
  for (int k = 0; k < nthreads; k++) pthread_create(&t[k], NULL, allpairs_worker, &J);
  for (int k = 0; k < nthreads; k++) pthread_join(t[k], NULL);
  free(t);
  return 0;
}
