#!/bin/bash
# Tuning sweep: library builds (min CTAs/SM) x columns per lane, C4 workload.
for lib in dafs_b200/libdafs_cuda.so exp/libdafs_mb5.so exp/libdafs_mb6.so exp/libdafs_mb8.so; do
  [ -f $lib ] || continue
  for c in 4 5 6 8 10; do
    echo -n "$lib cols=$c: "
    DAFS_CUDA_LIB=$PWD/$lib DAFS_PAIRHMM_COLS=$c python tools/prof_pairhmm.py ${1:-256} ${2:-300} 3 | sed 's/.*ms \([0-9.]*\) .*/\1 ms/'
  done
done
