/* libdafs_cuda -- C ABI of the B200-native DAFS hot path.
 *
 * Drop-in boundary for satoken/dafs (reference paths are relative to its tree):
 *   (1) all-pairs ProbCons pair-HMM matching posteriors
 *         replaces Align::Model::calculate(fa, mp)        src/align.cpp:35-52
 *         and ProbCons::calculate(seq1, seq2, mp)         src/align.cpp:60-79
 *   (2) Boltzmann/McCaskill base-pairing probabilities
 *         replaces Fold::Model::calculate(fa, bp)         src/fold.cpp:60-68
 *         and RNAfold::calculate(seq, bp)                 src/fold.cpp:78-116
 *   (3) the dual-decomposition solver and its two decoders
 *         replaces DAFS::solve_by_dd                      src/dafs.cpp:1006-1295
 *         SparseNussinov::decode                          src/nussinov.cpp:207-392
 *         SparseNeedlemanWunsch::initialize/decode        src/needleman_wunsch.cpp:198-422
 *   (4) the consistency transformation of DAFS::run()     src/dafs.cpp:1796-1827
 *   (5) the progressive alignment along the guide tree
 *         replaces DAFS::build_tree                       src/dafs.cpp:447-495
 *         and DAFS::align(ss, aln, root)                  src/dafs.cpp:1518-1537
 *
 * Plain C types only.  Every entry point returns 0 on success or a negative
 * DAFS_CUDA_E* code; dafs_cuda_last_error() gives the message (thread-local).
 * No exception crosses this boundary.  There is no CPU fallback: without a
 * usable CUDA device every compute entry point fails with DAFS_CUDA_ENODEV.
 */
#ifndef DAFS_CUDA_H_
#define DAFS_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DAFS_CUDA_OK 0
#define DAFS_CUDA_EINVAL (-1)   /* bad argument                               */
#define DAFS_CUDA_ENODEV (-2)   /* no CUDA device / device index out of range */
#define DAFS_CUDA_ECUDA (-3)    /* CUDA runtime error                         */
#define DAFS_CUDA_ENOMEM (-4)   /* host or device allocation failed           */
#define DAFS_CUDA_ENOTSUP (-5)  /* not supported (like the reference's throws) */

/* ---- lifecycle (no reference equivalent; errors there are C++ throws caught
 *      in main, src/dafs.cpp:1898-1909) -------------------------------------- */
int dafs_cuda_init(int ngpu /* 0 = all visible */);
void dafs_cuda_shutdown(void);
const char* dafs_cuda_last_error(void);
int dafs_cuda_device_count(void);
/* Measured FP32 issue rate of the current device, the denominator of the pair-HMM
 * roofline (MEASURED_PEAKS.json has no FP32 entry): dependent-chain
 * FFMA (2 flop each) and FMUL+FADD (the unfused arithmetic bit-exact parity
 * allows).  TFLOP/s, best of 3 after a warm-up launch.                          */
int dafs_cuda_fp32_probe(double* fma_tflops, double* muladd_tflops);
/* Same for FP64 (DFMA): the denominator of the McCaskill stage's roofline.        */
int dafs_cuda_fp64_probe(double* fma_tflops);

/* ---- sparse results ------------------------------------------------------
 * A batch of row-sparse matrices, one per unit (a sequence pair for (1), a
 * sequence for (2)), in the order the units were given.  It carries exactly
 * what the reference keeps in MP / BP = vector<vector<pair<uint,float>>>
 * (src/typedefs.h:39-41): for unit u and row r the kept entries are
 *   k in [ row_ptr[row_off[u] + r], row_ptr[row_off[u] + r + 1] )  (+ nz_off[u])
 * with col[] ascending inside a row (src/dafs.cpp:533-544 relies on that) and
 * val[] strictly greater than the threshold.  All arrays are owned by the
 * library (pinned host memory) until dafs_cuda_free_batch().                */
typedef struct dafs_csr_batch {
  int64_t n_units;
  int64_t nnz;            /* total kept entries                                   */
  const int64_t* row_off; /* [n_units+1] start of unit u's row_ptr block           */
  const int32_t* row_ptr; /* per unit: rows+1 entries, relative to nz_off[u]       */
  const int64_t* nz_off;  /* [n_units+1] start of unit u's entries in col/val      */
  const uint32_t* col;    /* [nnz] 0-based column                                  */
  const float* val;       /* [nnz]                                                 */
  void* owner_;           /* private                                               */
} dafs_csr_batch;

void dafs_cuda_free_batch(dafs_csr_batch* b);

/* ---- (1) pair-HMM posteriors ---------------------------------------------
 * seq[i] are the raw sequence bytes (no terminator needed), len[i] >= 1.
 * Alphabet handling equals the reference's (case-insensitive ACGUTN, anything
 * else = its "unknown" emission, probconsRNA/wrapper.cpp:134-171).
 * th is Align::Model::threshold() (-u, default 0.01).                       */

/* All i<j pairs in row-major order: unit index = pair rank in
 *   for i in [0,nseq) for j in (i,nseq)            (src/align.cpp:41-51).
 * ngpu: 0/1 = current device only; k>1 = shard pairs over devices 0..k-1 of
 * this process and merge on the host.                                       */
int dafs_cuda_pairhmm_all(int nseq, const char* const* seq, const int* len, float th,
                          int ngpu, dafs_csr_batch* out);

/* An explicit list of (pi[k], pj[k]) pairs; rows come from seq[pi], columns
 * from seq[pj].  This is what one rank calls when pairs are sharded across
 * processes (one process per GPU).                                          */
int dafs_cuda_pairhmm_pairs(int nseq, const char* const* seq, const int* len,
                            int64_t npairs, const int32_t* pi, const int32_t* pj,
                            float th, dafs_csr_batch* out);

/* Split-phase form of the same call, for callers that keep inputs resident in
 * HBM (and for timing the device part alone):
 *   create  = encode + H2D of sequences and pair list, workspace allocation
 *   run     = all kernels, asynchronous on the plan's stream; no host copies
 *   sync    = wait; optionally reports the device time of the last run (ms,
 *             CUDA events on the plan's stream)
 *   fetch   = D2H of the CSR result into a batch (implies sync)
 * A plan belongs to the device that was current at create.                  */
typedef struct dafs_pairhmm_plan dafs_pairhmm_plan;
int dafs_cuda_pairhmm_plan_create(int nseq, const char* const* seq, const int* len,
                                  int64_t npairs, const int32_t* pi, const int32_t* pj,
                                  float th, dafs_pairhmm_plan** plan);
int dafs_cuda_pairhmm_plan_run(dafs_pairhmm_plan* plan);
int dafs_cuda_pairhmm_plan_sync(dafs_pairhmm_plan* plan, float* device_ms, int64_t* nnz);
int dafs_cuda_pairhmm_plan_fetch(dafs_pairhmm_plan* plan, dafs_csr_batch* out);
void dafs_cuda_pairhmm_plan_destroy(dafs_pairhmm_plan* plan);

/* The same result delivered into HOST arrays the caller owns -- what one rank of a
 * multi-process run (one process per GPU, pairs sharded with dafs_cuda_pairhmm_pairs'
 * explicit list) calls with its region of a host buffer that the consuming process maps,
 * e.g. POSIX shared memory registered with cudaHostRegister; replaces the per-pair
 * push_back into mp[i][j] of src/align.cpp:69-78 for that rank's pairs.  The caller sets
 * the cap_* fields and the four pointers; the call fills n_units, rows, nnz and the arrays
 * (same meaning as in dafs_csr_batch; the row_ptr block of unit u starts at the sum of
 * len[pi[k]] + 1 over k < u).  Too small: DAFS_CUDA_ENOMEM with n_units / rows / nnz set
 * to what is needed.  Page-locked arrays let the copy of a finished part of the pair list
 * run under the kernel of the next part.                                              */
typedef struct dafs_host_csr {
  int64_t cap_units, cap_rows, cap_nnz;   /* in: capacity of nz_off (- 1), row_ptr, col / val */
  int64_t* nz_off;                        /* [cap_units + 1]                                   */
  int32_t* row_ptr;                       /* [cap_rows]                                        */
  uint32_t* col;                          /* [cap_nnz]                                         */
  float* val;                             /* [cap_nnz]                                         */
  int64_t n_units, rows, nnz;             /* out                                               */
} dafs_host_csr;
int dafs_cuda_pairhmm_pairs_into(int nseq, const char* const* seq, const int* len,
                                 int64_t npairs, const int32_t* pi, const int32_t* pj,
                                 float th, dafs_host_csr* dst);
int dafs_cuda_pairhmm_plan_fetch_into(dafs_pairhmm_plan* plan, dafs_host_csr* dst);

/* Device-resident batches.  fetch_device returns a dafs_csr_batch whose row_off / nz_off are host
 * arrays and whose row_ptr / col / val are DEVICE pointers (the plan's own arrays: the plan gives them
 * up and cannot be run again); dafs_cuda_pct, dafs_cuda_pct_device and dafs_cuda_progressive accept
 * such batches wherever they accept host ones, so the posteriors stay in HBM between the stages of
 * DAFS::run() (src/dafs.cpp:1787-1838) instead of crossing PCIe twice.  dafs_cuda_free_batch releases
 * either kind; dafs_cuda_batch_to_host makes a host copy (tests, output).                          */
int dafs_cuda_pairhmm_plan_fetch_device(dafs_pairhmm_plan* plan, dafs_csr_batch* out);
int dafs_cuda_batch_is_device(const dafs_csr_batch* b);
int dafs_cuda_batch_to_host(const dafs_csr_batch* b, dafs_csr_batch* out);

/* Device-resident hand-over: copies the pair-ordered CSR (row_ptr blocks,
 * nz_off[npairs+1], col, val) into caller-owned DEVICE buffers, e.g. the
 * tensors a rank feeds to an NCCL all-gather.  cap_nnz = capacity of d_col/d_val
 * in entries; any pointer may be NULL to skip that array.  Implies sync.     */
int dafs_cuda_pairhmm_plan_export(dafs_pairhmm_plan* plan, int32_t* d_row_ptr,
                                  int64_t* d_nz_off, uint32_t* d_col, float* d_val,
                                  int64_t cap_nnz);

/* ---- K6: hand-over between ranks (one process per GPU, NCCL over NVLink) ---------
 * No counterpart in the single-process reference; this is the "sparse posterior matrices
 * are gathered to the rank that runs progressive alignment" step, i.e.
 * what stands between the per-rank loops over pairs (src/align.cpp:41-51) and the code
 * that reads all of mp_ (src/dafs.cpp:1797-1838).  NCCL is bound at run time
 * (libnccl.so.2); without it these calls return DAFS_CUDA_ENOTSUP.
 *   unique_id  rank 0 makes the 128-byte id and hands it to the others (any channel)
 *   create     collective over all ranks; binds the communicator to the current device
 *   gather     collective: every rank's fragment (the result of its plan, pairs in that
 *              rank's own order) lands on `root`, device-resident, exact sizes, as the
 *              concatenation over ranks r of
 *                d_nz_off  [units_r + 1] entries at unit_off[r] + r, relative to nnz_off[r]
 *                d_row_ptr [rows_r] at rows_off[r];  d_col / d_val [nnz_r] at nnz_off[r]
 *              (unit_off / nnz_off / rows_off are host arrays of nranks + 1 prefix sums).
 *              On the other ranks *out stays empty.                                     */
#define DAFS_CUDA_COMM_ID_BYTES 128
typedef struct dafs_comm dafs_comm;
int dafs_cuda_comm_unique_id(void* id128);
int dafs_cuda_comm_create(const void* id128, int rank, int nranks, dafs_comm** comm);
int dafs_cuda_comm_info(const dafs_comm* comm, int* rank, int* nranks, int* nccl_version);
void dafs_cuda_comm_destroy(dafs_comm* comm);
typedef struct dafs_gathered_csr {
  int32_t nranks;
  const int64_t* unit_off;   /* host [nranks+1] */
  const int64_t* nnz_off;    /* host [nranks+1] */
  const int64_t* rows_off;   /* host [nranks+1] */
  const int64_t* d_nz_off;   /* device */
  const int32_t* d_row_ptr;  /* device */
  const uint32_t* d_col;     /* device */
  const float* d_val;        /* device */
  void* owner_;              /* private */
} dafs_gathered_csr;
int dafs_cuda_pairhmm_plan_gather(dafs_pairhmm_plan* plan, dafs_comm* comm, int root,
                                  dafs_gathered_csr* out);
/* D2H of a gathered result into caller arrays sized from the offsets (any may be NULL). */
int dafs_cuda_gathered_to_host(const dafs_gathered_csr* g, int64_t* nz_off, int32_t* row_ptr,
                               uint32_t* col, float* val);
void dafs_cuda_free_gathered(dafs_gathered_csr* g);

/* Geometry and timings of a plan (for benchmarks and profiles). */
typedef struct dafs_pairhmm_info {
  int cols_per_lane;        /* C: DP columns each lane keeps in registers          */
  int grid;                 /* persistent CTAs                                     */
  int warps_per_cta;
  int launches_per_run;     /* kernels one run() launches                          */
  int64_t workspace_bytes;  /* per-warp F_M planes in HBM                          */
  float last_total_ms;      /* device time of the last run, all kernels            */
  float last_kernel_ms;     /* ... of the forward/backward/posterior kernel alone  */
} dafs_pairhmm_info;
int dafs_cuda_pairhmm_plan_info(dafs_pairhmm_plan* plan, dafs_pairhmm_info* info);

/* Log-space model tables the kernels use (3 init, 9 trans, 6 insert, 36 match;
 * symbol classes A C G U/T N other) -- exported so tests can pin them against
 * the reference's ProbabilisticModel constructor (ProbabilisticModel.h:55-88). */
int dafs_cuda_pairhmm_tables(float* out54);

/* ---- (2) McCaskill base-pairing probabilities -----------------------------
 * What RNAfold(bl = true)::calculate(seq, bp) delivers (src/fold.cpp:78-116:
 * ViennaRNA pf_fold + export_bppm under the BL* tables of
 * src/boltzmann_param.c): per sequence the upper-triangular rows i < j with
 * float(p) > th, columns ascending.  th is Fold::Model::threshold() (CUTOFF =
 * 0.01, src/dafs.cpp:65,1700).  Alphabet: ACGU, T = U, case-insensitive, other
 * bytes cannot pair (ViennaRNA's encoding).  1 <= len[i] <= 6000.
 * Unit u of the batch = sequence u, len[u]+1 row_ptr entries.                */
int dafs_cuda_mccaskill_all(int nseq, const char* const* seq, const int* len, float th, int ngpu,
                            dafs_csr_batch* out);

/* Split-phase form, like the pair-HMM plan.  fetch: lnz may be NULL or receive
 * ln Z per sequence (= -ensemble free energy / kT).                          */
typedef struct dafs_mccaskill_plan dafs_mccaskill_plan;
int dafs_cuda_mccaskill_plan_create(int nseq, const char* const* seq, const int* len, float th,
                                    dafs_mccaskill_plan** plan);
/* Constrained folding: RNAfold::calculate(seq, str, bp), src/fold.cpp:118-166 (what
 * --bp-update / --bp-update1 / --ipknot reach through update_basepairing_probability,
 * src/dafs.cpp:609-711).  str[i] is the constraint of sequence i in the reference's alphabet
 * ('?' no constraint, '.' stays unpaired, '(' ')' a given pair; the reference rewrites it to
 * ViennaRNA's '.', 'x', '(' ')' at :124-125), as long as the sequence; str == NULL or
 * str[i] == NULL = unconstrained.  ViennaRNA's pf_fold applies such a constraint by removing
 * the conflicting pairs: a '.' base pairs with nobody, the ends of a given pair pair with
 * nobody else, and no pair crosses a given pair.                                        */
int dafs_cuda_mccaskill_plan_create_constrained(int nseq, const char* const* seq, const int* len,
                                                const char* const* str, float th,
                                                dafs_mccaskill_plan** plan);
int dafs_cuda_mccaskill_constrained(int nseq, const char* const* seq, const int* len,
                                    const char* const* str, float th, dafs_csr_batch* out);
int dafs_cuda_mccaskill_plan_run(dafs_mccaskill_plan* plan);
int dafs_cuda_mccaskill_plan_sync(dafs_mccaskill_plan* plan, float* device_ms, int64_t* nnz);
int dafs_cuda_mccaskill_plan_fetch(dafs_mccaskill_plan* plan, dafs_csr_batch* out, double* lnz);
int dafs_cuda_mccaskill_plan_fetch_device(dafs_mccaskill_plan* plan, dafs_csr_batch* out);
void dafs_cuda_mccaskill_plan_destroy(dafs_mccaskill_plan* plan);
typedef struct dafs_mccaskill_info {
  int grid, block, smem_bytes, row_stride, table_bytes;
  int64_t workspace_bytes;
  float last_ms;
} dafs_mccaskill_info;
int dafs_cuda_mccaskill_plan_info(dafs_mccaskill_plan* plan, dafs_mccaskill_info* info);

/* ---- (3) dual decomposition ------------------------------------------------
 * One problem = one call of DAFS::solve_by_dd(x, y, z, p_x, p_y, p_z, aln1, aln2)
 * (src/dafs.cpp:1006): dense row-major probability matrices exactly as the
 * reference passes them (VVF), N1 = aln1.size(), N2 = aln2.size() (:1013-1014),
 * and the members it reads: w = w_, th_s = th_s_[0] (the threshold of
 * s_decoder_, :1759), min_th_s = min(th_s_) (:1018), th_a = th_a_, eta0 = eta0_,
 * t_max = t_max_ (-m).  Requires 1 <= L1, L2 <= 16384.                      */
typedef struct dafs_dd_problem {
  int32_t L1, L2, N1, N2;
  const float* p_x;   /* [L1*L1], only i<j read */
  const float* p_y;   /* [L2*L2]                */
  const float* p_z;   /* [L1*L2]                */
  float w, th_s, min_th_s, th_a, eta0;
  int32_t t_max;
} dafs_dd_problem;

/* x, y, z are caller-allocated (L1, L2, L1 entries; 0xffffffff = "-1u", no
 * partner), trace may be NULL or hold 3*t_max floats {L, eta, violated} per
 * iteration (what -v 2 logs, src/dafs.cpp:1273).  score is solve_by_dd's return
 * value, iters / violated what it logs at :1292.  status != 0: the NW traceback
 * met an unfilled cell (the reference would not terminate there).           */
typedef struct dafs_dd_result {
  uint32_t* x;
  uint32_t* y;
  uint32_t* z;
  float* trace;
  float score;
  int32_t iters;
  uint32_t violated;
  int32_t status;
  int64_t ncbp;       /* consensus-base-pair candidates enumerated (:1016-1045) */
  float cycles[8];    /* SM cycles: x fill, x traceback, y fill, y traceback, z fill,
                         z traceback, decode phase, tally+update phase             */
} dafs_dd_result;

/* All problems in one launch, each running its whole <= t_max iteration loop on
 * the device (no host round trip between iterations).                       */
int dafs_cuda_dd_solve_batch(int nprob, const dafs_dd_problem* prob, dafs_dd_result* res);

/* Split-phase form (inputs resident in HBM; run() is repeatable):
 *   create = CBP enumeration (:1016-1061) + NW envelope (needleman_wunsch.cpp:
 *            198-253) + H2D;   run = reset multipliers + the DD kernel;
 *   sync   = wait, device ms;  fetch = D2H of x, y, z, score, iters, violated. */
typedef struct dafs_dd_plan dafs_dd_plan;
int dafs_cuda_dd_plan_create(int nprob, const dafs_dd_problem* prob, int want_trace,
                             dafs_dd_plan** plan);
/* The same plan from matrices that are already in HBM: p_x / p_y / p_z of every problem are DEVICE
 * pointers (e.g. the averaged matrices of guide-tree nodes, which dafs_cuda_progressive builds on
 * the device).  CBP enumeration (:1016-1061), the NW envelope and all packing run in two kernels;
 * the host sees 64 bytes of counts per problem.  The matrices may be reused once the call returns.
 * The resulting plan is byte-identical to the one dafs_cuda_dd_plan_create builds.            */
int dafs_cuda_dd_plan_create_device(int nprob, const dafs_dd_problem* prob, int want_trace,
                                    dafs_dd_plan** plan);
int dafs_cuda_dd_plan_run(dafs_dd_plan* plan);
int dafs_cuda_dd_plan_sync(dafs_dd_plan* plan, float* device_ms);
int dafs_cuda_dd_plan_fetch(dafs_dd_plan* plan, dafs_dd_result* res);
/* Non-blocking harvest while run() is in flight: every problem that has finished and whose
 * done[k] is still 0 is written to res[k] (same fields as fetch) and done[k] set to 1;
 * *n_new = how many.  The caller keeps done[] (nprob entries, zeroed before run).  This is what
 * lets dafs_cuda_progressive start a guide-tree node as soon as its two children are solved. */
int dafs_cuda_dd_plan_poll(dafs_dd_plan* plan, dafs_dd_result* res, int32_t* done, int32_t* n_new);
void dafs_cuda_dd_plan_destroy(dafs_dd_plan* plan);

typedef struct dafs_dd_info {
  int grid, block, cols_per_lane;  /* persistent CTAs, threads, 32-column words of the widest NW band */
  int smem_bytes;                  /* dynamic shared memory per CTA                 */
  int ctas_per_sm;
  unsigned items_in_smem;          /* bit per scratch array (dd_kernel.cuh)         */
  int64_t workspace_bytes;         /* per-CTA global scratch, all CTAs              */
  int64_t arena_bytes;             /* per-problem device data                       */
  int64_t total_cbp, total_entries;
  float last_ms;
  int packed_on_device;            /* 1: the constant blocks were built by the pack kernels */
} dafs_dd_info;
int dafs_cuda_dd_plan_info(dafs_dd_plan* plan, dafs_dd_info* info);

/* Single decodes, for the reference's other call sites of the two decoders
 * (src/dafs.cpp:923, :1867, :1870) and for bit-exactness tests.
 * Nussinov: p, q are [L*L]; q == NULL selects the plain overload
 * (src/nussinov.cpp:300-392: s = p - th).  NW: p, q are [L1*L2]; includes
 * initialize() (src/needleman_wunsch.cpp:198-253); q == NULL = plain overload. */
int dafs_cuda_nussinov_decode(int L, float th, float w, const float* p, const float* q,
                              uint32_t* ss, float* score);
int dafs_cuda_nw_decode(int L1, int L2, float th, const float* p, const float* q, uint32_t* al,
                        float* score);

/* Allocation-free variant of dafs_cuda_nw_decode.  It has no counterpart in the
 * reference; it was added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same results (bit for bit).  All
 * device memory it uses -- band image of p / q, envelope, DP rows, traceback
 * planes, result -- is one file-scope device array of DAFS_CUDA_NW_STATIC_BYTES
 * that every GPU loading the library holds until exit; the host image of the
 * inputs is a static host array of the same size.  It calls no cudaMalloc,
 * cudaFree, cudaHostAlloc or other allocating routine and uses no host heap.
 * Limits: L1, L2 <= DAFS_CUDA_NW_STATIC_MAXL and a scratch need (about 8 bytes per
 * envelope cell + 16 * (L1 + L2) + (L1 + 1) * ceil(widest band / 32) * 8) within the
 * array; anything larger is rejected with DAFS_CUDA_EINVAL and a message naming the
 * limit.  The array is shared by all calls and keeps the last call's contents
 * (what must start at zero is re-uploaded by each call); calls are serialised by a
 * host mutex and the legacy default stream -- calls that overlap on the GPU race. */
#define DAFS_CUDA_NW_STATIC_BYTES (8 << 20)
#define DAFS_CUDA_NW_STATIC_MAXL 2048
int dafs_cuda_nw_decode_static(int L1, int L2, float th, const float* p, const float* q,
                               uint32_t* al, float* score);

/* ---- (4) probabilistic consistency transformation --------------------------
 * The stage DAFS::run() executes between the posteriors and the guide tree
 * (src/dafs.cpp:1796-1827): transpose_mp for every pair (:155-167, :1797-1799),
 * calculate_similarity_score for every pair (:713-764, :1812-1819),
 * relax_basepairing_probability when w_pct_s != 0 (:326-375, option -q) and
 * relax_matching_probability when w_pct_a != 0 (:258-324, option -p), in that
 * order; the defaults are 0.25 / 0.25 (:1622, :1631).  Bit-exact with the
 * reference: every cell is accumulated in the reference's (z, k) / (y, k, l)
 * order with unfused FP32 arithmetic and its mixed float / double weights.
 *   mp      all i<j pairs, row-major, rows = len[i]  (what dafs_cuda_pairhmm_all returns)
 *   bp      one unit per sequence, rows = len        (what dafs_cuda_mccaskill_all returns)
 *   sim     out, [nseq*nseq] row-major, or NULL
 *   mp_out  relaxed matching probabilities, same layout as mp (the reference's
 *           mp_[x][y] for x < y; mp_[y][x] is its transpose, mp_[x][x] the identity)
 *   bp_out  relaxed base-pairing probabilities, same layout as bp
 * With a weight of 0 the corresponding output is a copy of the input.           */
int dafs_cuda_pct(int nseq, const int* len, const dafs_csr_batch* mp, const dafs_csr_batch* bp,
                  float w_pct_a, float w_pct_s, float* sim, dafs_csr_batch* mp_out,
                  dafs_csr_batch* bp_out);

/* The same with device-resident outputs (see dafs_cuda_pairhmm_plan_fetch_device). */
int dafs_cuda_pct_device(int nseq, const int* len, const dafs_csr_batch* mp, const dafs_csr_batch* bp,
                         float w_pct_a, float w_pct_s, float* sim, dafs_csr_batch* mp_out,
                         dafs_csr_batch* bp_out);

/* Device time of the kernels of the last dafs_cuda_pct call on this thread, in ms:
 * {transpose, similarity, relax_bp, relax_mp, total incl. H2D/D2H}.              */
void dafs_cuda_pct_last_ms(float out5[5]);

/* ---- (5) progressive alignment along the guide tree ---------------------------
 * dafs_cuda_guide_tree: DAFS::build_tree (src/dafs.cpp:447-495) from the similarity
 * matrix sim[nseq*nseq] (what dafs_cuda_pct returns).  Nodes 0..nseq-1 are the leaves
 * (left = right = -1), nodes nseq..2*nseq-2 the merges in the order they were made, the
 * last one the root; score[k] = tree_[k].first (may be NULL).  Host arithmetic only.   */
int dafs_cuda_guide_tree(int nseq, const float* sim, int32_t* left, int32_t* right, float* score);

/* dafs_cuda_progressive: DAFS::align(ss, aln, root) (src/dafs.cpp:1518-1537), i.e. for
 * every internal node align_alignments (:913-951): average_basepairing_probability of both
 * children (:557-607), average_matching_probability (:513-555), solve_by_dd (:1006-1295),
 * project_alignment (:766-826); for the root also project_secondary_structure (:828-876)
 * and the consensus (:941-945).  Nodes whose children are finished are processed together
 * (one averaging launch pair + one batched DD launch per batch of ready nodes, harvested
 * problem by problem while the launch is still running); sibling subtrees share no
 * state, so the results equal the reference's depth-first recursion bit for bit.
 * Covers use_alifold_ == false (--no-alifold; RNAalifold needs ViennaRNA) and
 * use_bp_update_ == false (the default).
 *   mp, bp   what DAFS::run() holds at :1829: all i<j pairs / one unit per sequence,
 *            e.g. the outputs of dafs_cuda_pct
 *   left/right  the guide tree, 2*nseq-1 nodes as dafs_cuda_guide_tree returns them
 *   prm      w = w_ (-w), th_s = th_s_[0] (-t), min_th_s = min(th_s_), th_a = th_a_ (-u),
 *            eta0 = eta0_ (--eta), t_max = t_max_ (-m)                                  */
typedef struct dafs_prog_params {
  float w, th_s, min_th_s, th_a, eta0;
  int32_t t_max;
} dafs_prog_params;

typedef struct dafs_prog_result {
  int32_t L;                 /* columns of the root alignment                            */
  int32_t n_rows;            /* = nseq                                                   */
  const int32_t* seq;        /* [n_rows] ALN row r is sequence seq[r] (the reference's row
                                order: rows of the left child, then of the right child)   */
  const uint8_t* bits;       /* [n_rows * L] ALN[r].second: 1 = residue, 0 = gap          */
  const uint32_t* ss;        /* [L] consensus structure of the root (partner column or
                                0xffffffff); NULL when nseq == 1                          */
  float score;               /* align()'s return value: the root's solve_by_dd score      */
  int32_t n_nodes;           /* 2*nseq-1; the arrays below are indexed by tree node       */
  const int32_t* node_iters; /* what solve_by_dd logs at :1292 (leaves: 0)                */
  const uint32_t* node_violated;
  const float* node_score;
  const int32_t* node_L;     /* alignment columns of the node                             */
  int32_t waves;             /* batched DD launches (a node goes out as soon as both its
                                children are solved; at least the height of the tree)      */
  int64_t total_iters;
  int32_t gpu_launches;
  float ms_average;          /* device time of the averaging kernels, all waves           */
  float ms_dd;               /* device time of the DD kernel, all waves                   */
  float ms_dd_host;          /* wall time of the DD plans incl. host set-up and copies    */
  float ms_total;            /* wall time of the call                                     */
  void* owner_;              /* private                                                   */
} dafs_prog_result;

int dafs_cuda_progressive(int nseq, const int* len, const dafs_csr_batch* mp, const dafs_csr_batch* bp,
                          const int32_t* left, const int32_t* right, const dafs_prog_params* prm,
                          dafs_prog_result* out);
void dafs_cuda_free_prog_result(dafs_prog_result* r);

#ifdef __cplusplus
}
#endif
#endif /* DAFS_CUDA_H_ */
