// TEST INFRASTRUCTURE ONLY (part of oracle/_ref/libdafs_ref.so, never shipped or measured).
// Reads the log-space tables out of the reference's own ProbabilisticModel constructor
// (src/probconsRNA/ProbabilisticModel.h:55-88) fed with the parameter globals that
// Probcons::Impl::ReadParameters fills (src/probconsRNA/wrapper.cpp:14-19, 134-171).
// The tables are private members: this one TU is compiled with -fno-access-control.
#include <string>
#include <vector>

#include "probconsRNA/ProbabilisticModel.h"
#include "probconsRNA/probcons.h"

namespace PROBCONS {
extern VF initDistrib, gapOpen, gapExtend, emitSingle;
extern VVF emitPairs;
}

// out: init[3] | trans[9] | ins[256] (state 0; the three states hold the same value, :83-84) |
//      match[256*256]
extern "C" void ref_probcons_tables(float* out) {
  PROBCONS::Probcons pc;   // its constructor runs ReadParameters()
  PROBCONS::ProbabilisticModel m(PROBCONS::initDistrib, PROBCONS::gapOpen, PROBCONS::gapExtend,
                                 PROBCONS::emitPairs, PROBCONS::emitSingle);
  int k = 0;
  for (int i = 0; i < 3; ++i) out[k++] = m.initialDistribution[i];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) out[k++] = m.transProb[i][j];
  for (int c = 0; c < 256; ++c) {
    out[k++] = m.insProb[c][0];
    if (m.insProb[c][1] != m.insProb[c][0] || m.insProb[c][2] != m.insProb[c][0]) out[0] = -1.0f / 0.0f;
  }
  for (int a = 0; a < 256; ++a)
    for (int b = 0; b < 256; ++b) out[k++] = m.matchProb[a][b];
}
