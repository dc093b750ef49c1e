#!/bin/bash
# One GPU-box session: tests, bench, ncu launch list, ncu captures of the round's new kernels.
# Everything lands in gpurun_out/ (prefix $1).  usage: tools/gpu_round.sh s14
P=${1:-run}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${P}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${P}_pytest.log
tail -3 $O/${P}_pytest.log
python bench.py > $O/${P}_bench.json 2> $O/${P}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/${P}_bench_ref.json 2> $O/${P}_bench_ref.err; echo "bench ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${P}_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 > $O/${P}_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:prog_avg_mp -c 1 -f -o $O/${P}_prog_avg_mp \
    python tools/bench_prog.py 256 300 random 600 0 1 > $O/${P}_ncu_avg_mp.log 2>&1; echo "ncu avg_mp rc=$?"
ncu --set full --clock-control none --import-source on -k regex:dd_kernel -c 1 -f -o $O/${P}_dd_rf17 \
    python tools/bench_prog.py 0 0 tests/golden/RF00017_4.fa 600 0 1 > $O/${P}_ncu_dd.log 2>&1; echo "ncu dd rc=$?"
ncu --set full --clock-control none --import-source on -k regex:pairhmm_kernel -s 1 -c 1 -f -o $O/${P}_pairhmm \
    python tools/prof_pairhmm.py 256 300 2 > $O/${P}_ncu_pairhmm.log 2>&1; echo "ncu pairhmm rc=$?"
