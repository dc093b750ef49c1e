// Energy model of (2): the "Boltzmann" fold model of the reference =
// ViennaRNA 2.4.18 pf_fold defaults (dangles 2, special hairpins, TURN 3,
// MAXLOOP 30, 37 C) over the BL* tables the reference copies into ViennaRNA's
// globals (src/boltzmann_param.c:6010-6026, used by RNAfold::RNAfold(bl=true),
// src/fold.cpp:70-76).  Boltzmann factors exp(-E/kT) are built once on the host
// in double, stored as FP32 and kept resident in shared memory by the kernel
// (62 KB); all partition-function arithmetic is FP64.
//
// Tables the reference leaves at ViennaRNA's Turner-2004 defaults are not
// available on this box; they are substituted exactly as DESIGN.md states:
// mismatchM / mismatchExt = dangle5 + dangle3, mismatch1nI = mismatch23I =
// mismatchI, no Triloops / Hexaloops, lxc37 = 107.856.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace dafs {

#define MC_TURN 3
#define MC_MAXLOOP 30

struct McTables {              // pair types 1..6 -> index 0..5; nucleotides 0 (other), 1..4 = A C G U
  float stack[6][6];
  float mmH[6][5][5], mmI[6][5][5], mmM[6][5][5], mmE[6][5][5];
  float d5[6][5], d3[6][5];
  float int11[6][6][5][5];
  float int21[6][6][5][5][5];
  float int22[6][6][4][4][4][4];   // nucleotides 1..4 only (what copy_int22 fills, boltzmann_param.c:5953-5963)
  float hairpin[31], bulge[31], internal[31], ninio[31];
  float ml_base, ml_closing, ml_intern, termAU;
  float lxc_over_kT;           // lxc37 * 10 / kT
  int ntetra;
  int tetra_code[32];          // six nucleotide codes, base 5
  float tetra[32];
};

// pair type of two nucleotide codes: CG 1, GC 2, GU 3, UG 4, AU 5, UA 6, else 0
__host__ __device__ __forceinline__ int mc_pair(int a, int b) {
  // packed 5x5 table, 3 bits each: rows a = 0..4
  // a=1 (A): b=4 -> 5 ; a=2 (C): b=3 -> 1 ; a=3 (G): b=2 -> 2, b=4 -> 3 ; a=4 (U): b=1 -> 6, b=3 -> 4
  const int code = a * 5 + b;
  return code == 9 ? 5 : code == 13 ? 1 : code == 17 ? 2 : code == 19 ? 3 : code == 21 ? 6 : code == 23 ? 4 : 0;
}
__host__ __device__ __forceinline__ int mc_rtype(int t) { return ((t - 1) ^ 1) + 1; }   // 1<->2, 3<->4, 5<->6

// exp_E_Hairpin.  s6 = base-5 code of S[i..i+5] (only used when u == 4).
__device__ __forceinline__ double mc_hairpin(const McTables& T, int u, int type, int si1, int sj1, int s6) {
  double q = (u <= 30) ? (double)T.hairpin[u] : (double)T.hairpin[30] * pow((double)u / 30.0, -(double)T.lxc_over_kT);
  if (u == 4) {
    for (int t = 0; t < T.ntetra; ++t)
      if (T.tetra_code[t] == s6) return (double)T.tetra[t];
  } else if (u == 3) {
    return type > 2 ? q * (double)T.termAU : q;
  }
  return q * (double)T.mmH[type - 1][si1][sj1];
}

// exp_E_IntLoop: type = closing pair (i,j), type2 = rtype of the inner pair (k,l),
// si1 = S[i+1], sj1 = S[j-1], sp1 = S[k-1], sq1 = S[l+1].
__device__ __forceinline__ double mc_interior(const McTables& T, int u1, int u2, int type, int type2, int si1,
                                              int sj1, int sp1, int sq1) {
  const int ul = max(u1, u2), us = min(u1, u2);
  const int a = type - 1, b = type2 - 1;
  if (ul == 0) return (double)T.stack[a][b];
  if (us == 0) {
    double z = (double)T.bulge[ul];
    if (ul == 1) z *= (double)T.stack[a][b];
    else { if (type > 2) z *= (double)T.termAU; if (type2 > 2) z *= (double)T.termAU; }
    return z;
  }
  if (us == 1) {
    if (ul == 1) return (double)T.int11[a][b][si1][sj1];
    if (ul == 2) return (u1 == 1) ? (double)T.int21[a][b][si1][sq1][sj1] : (double)T.int21[b][a][sq1][si1][sp1];
  } else if (us == 2 && ul == 2) {
    if (si1 && sp1 && sq1 && sj1) return (double)T.int22[a][b][si1 - 1][sp1 - 1][sq1 - 1][sj1 - 1];
    return 1.0;   // a non-ACGU neighbour: the reference's table entry is ViennaRNA's default, unknown here
  }
  return (double)T.internal[ul + us] * (double)T.mmI[a][si1][sj1] * (double)T.mmI[b][sq1][sp1] * (double)T.ninio[ul - us];
}

// exp_E_MLstem (multi = true) / exp_E_ExtLoop with dangles = 2; si1 / sj1 < 0 beyond the ends.
__device__ __forceinline__ double mc_stem(const McTables& T, int type, int si1, int sj1, bool multi) {
  double e = 1.0;
  if (si1 >= 0 && sj1 >= 0) e = multi ? (double)T.mmM[type - 1][si1][sj1] : (double)T.mmE[type - 1][si1][sj1];
  else if (si1 >= 0) e = (double)T.d5[type - 1][si1];
  else if (sj1 >= 0) e = (double)T.d3[type - 1][sj1];
  if (type > 2) e *= (double)T.termAU;
  if (multi) e *= (double)T.ml_intern;
  return e;
}

}  // namespace dafs
