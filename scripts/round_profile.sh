#!/bin/bash
# One gpurun call: GPU tests, smoke, the bench lines (default workload + two fixtures + the CPU arm), arithmetic and
# fixture timings.  Every step runs under its own `timeout`.  Everything lands in gpurun_out/; the summaries that are
# kept are copied to profiles/ by hand (scripts/summarise_profiles.py for the ncu exports of round_ncu.sh).
set -u
O=gpurun_out; mkdir -p $O
(time timeout 1700 python -m pytest tests -m gpu -x -q) > $O/gpu_tests.log 2>&1; tail -3 $O/gpu_tests.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; cut -c1-200 $O/bench.json
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_reference.json 2> $O/bench_reference.err; cut -c1-200 $O/bench_reference.json
for w in burkardt_adlittle netlib_SC205 netlib_ISRAEL miplib_50v-10; do
  timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-k1-dense > $O/bench_$w.json 2> $O/bench_$w.err; cut -c1-200 $O/bench_$w.json
done
timeout 600 python scripts/arith_time.py $O/arith_time.jsonl > $O/arith_time.log 2>&1; tail -3 $O/arith_time.log
rm -f $O/netlib_times.jsonl; timeout 1200 python scripts/netlib_time.py > $O/netlib_time.log 2>&1; tail -2 $O/netlib_time.log
ls -la $O
