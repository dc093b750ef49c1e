#!/bin/sh
# Pins path (2) -- the "Boltzmann" fold model -- to the reference's own arithmetic.  UNTESTED HERE:
# neither the development container nor the B200 boxes have ViennaRNA, so
# this is the recipe for a maintainer on a machine with network access.  It
#   1. builds ViennaRNA 2.4.18 exactly as the reference's Dockerfile does (Dockerfile:21),
#   2. compiles a small driver against the reference's own fold.cpp + boltzmann_param.c and RNAlib2,
#   3. writes tests/golden/mccaskill_golden.json: for RF00005, RF00017 and three synthetic 300-mers
#      the (i, j, p) triples that RNAfold(bl = true)::calculate returns (src/fold.cpp:78-116) and,
#      for two constraint strings each, what the constrained overload returns (src/fold.cpp:118-166),
#   4. dumps the energy tables the BL* copy does NOT overwrite (mismatchM37, mismatchExt37,
#      mismatch1nI37, mismatch23I37, Triloops, Hexaloops, lxc37: src/boltzmann_param.c:6010-6026)
#      to tests/golden/vienna_defaults.json, which dafs_b200/csrc/mccaskill.cu substitutes today.
# tests/test_mccaskill_gpu.py::test_vienna_golden (skipped while the file is absent) then compares the
# device result with it at 1e-4.
#   usage: tools/pin_mccaskill.sh /path/to/dafs-reference [/path/to/prefix]
set -e
REF=${1:?path to the satoken/dafs checkout}
PREFIX=${2:-$PWD/_vienna}
HERE=$(cd "$(dirname "$0")/.." && pwd)
if ! pkg-config --exists RNAlib2; then
  mkdir -p "$PREFIX/src" && cd "$PREFIX/src"
  wget -q https://www.tbi.univie.ac.at/RNA/download/sourcecode/2_4_x/ViennaRNA-2.4.18.tar.gz
  tar zxf ViennaRNA-2.4.18.tar.gz && cd ViennaRNA-2.4.18
  ./configure --prefix="$PREFIX" --without-perl --without-python --without-python3 --without-forester --without-rnalocmin
  make -j"$(nproc)" && make install
  export PKG_CONFIG_PATH="$PREFIX/lib/pkgconfig:$PKG_CONFIG_PATH"
fi
cd "$HERE"
cat > /tmp/pin_mccaskill.cpp <<'CPP'
// driver: the reference's own RNAfold model on a FASTA file -> JSON lines
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>
#include "fold.h"
int main(int argc, char** argv) {
  RNAfold model(true, NULL, 0.01f);
  for (int a = 1; a < argc; ++a) {
    std::ifstream in(argv[a]);
    std::string line, name, seq;
    std::vector<std::pair<std::string, std::string> > fa;
    while (std::getline(in, line)) {
      if (!line.empty() && line[0] == '>') { if (!seq.empty()) fa.push_back({name, seq}); name = line.substr(1); seq.clear(); }
      else seq += line;
    }
    if (!seq.empty()) fa.push_back({name, seq});
    for (auto& f : fa) {
      for (int mode = 0; mode < 3; ++mode) {
        std::string con(f.second.size(), '?');
        if (mode == 1) for (size_t k = 0; k < con.size(); k += 7) con[k] = '.';
        if (mode == 2 && con.size() > 20) { con[2] = '('; con[con.size() - 3] = ')'; }
        BP bp;
        if (mode == 0) model.calculate(f.second, bp); else model.calculate(f.second, con, bp);
        std::printf("{\"seq\": \"%s\", \"con\": %s%s%s, \"bp\": [", f.second.c_str(), mode ? "\"" : "", mode ? con.c_str() : "null", mode ? "\"" : "");
        bool first = true;
        for (size_t i = 0; i < bp.size(); ++i)
          for (auto& e : bp[i]) { std::printf("%s[%zu, %u, %.9g]", first ? "" : ", ", i, e.first, e.second); first = false; }
        std::printf("]}\n");
      }
    }
  }
  return 0;
}
CPP
g++ -std=c++17 -O2 -DHAVE_VIENNA20 -I"$REF/src" $(pkg-config --cflags RNAlib2) -o /tmp/pin_mccaskill \
    /tmp/pin_mccaskill.cpp "$REF/src/fold.cpp" "$REF/src/boltzmann_param.c" "$REF/src/fa.cpp" \
    $(pkg-config --libs RNAlib2) -lm
python3 - <<'PY'
import sys; sys.path.insert(0, ".")
from dafs_b200.synth import random_rna
open("/tmp/pin_random.fa", "w").write("".join(">r%d\n%s\n" % (k, s.decode()) for k, s in enumerate(random_rna(3, 300, config_id=4))))
PY
/tmp/pin_mccaskill tests/golden/RF00005_0.fa tests/golden/RF00017_4.fa /tmp/pin_random.fa > /tmp/pin.jsonl
python3 - <<'PY'
import json
cases = [json.loads(l) for l in open("/tmp/pin.jsonl")]
json.dump({"source": "RNAfold(bl=true)::calculate of satoken/dafs linked to ViennaRNA 2.4.18", "cases": cases},
          open("tests/golden/mccaskill_golden.json", "w"))
print("wrote tests/golden/mccaskill_golden.json:", len(cases), "cases")
PY
echo "now run: python -m pytest tests/test_mccaskill_gpu.py -k vienna_golden -m gpu"
