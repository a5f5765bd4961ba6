#!/bin/bash
# One gpurun call: GPU tests, the bench line, ncu launch list + full captures (each after the same command exited 0
# without ncu), netlib timings.  Everything lands in gpurun_out/; scripts/summarise_profiles.py turns it into profiles/.
set -u
O=gpurun_out; mkdir -p $O
(time python -m pytest tests -m gpu -x -q) > $O/gpu_tests.log 2>&1; tail -3 $O/gpu_tests.log
python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; cut -c1-160 $O/bench.json
python bench.py --steps 2 --warmup 3 --cpu-pivots 10 --no-k1-dense > $O/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 3 --cpu-pivots 10 --no-k1-dense > $O/ncu_bench.log 2>&1
python bench.py --steps 1 --warmup 3 --cpu-pivots 10 --no-k1-dense > $O/plain_bench1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_simplex_loop --launch-skip 7 --launch-count 1 -o $O/loop \
    python bench.py --steps 1 --warmup 3 --cpu-pivots 10 --no-k1-dense > $O/ncu_loop.log 2>&1
python scripts/k1_dense.py 20000 3 > $O/k1_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_update$ -s 1 -c 1 -o $O/k1_dense \
    python scripts/k1_dense.py 20000 3 > $O/k1_ncu.log 2>&1
python scripts/k1_dense.py 20000 8 | tail -1 > $O/k1_dense.json; cat $O/k1_dense.json
rm -f $O/netlib_times.jsonl; python scripts/netlib_time.py > $O/netlib_time.log 2>&1; tail -2 $O/netlib_time.log
ls -la $O
