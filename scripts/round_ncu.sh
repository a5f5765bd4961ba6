#!/bin/bash
# The round's ncu passes (one GPU; each capture only after the same command exited 0 without ncu): launch list of the
# bench command, --set full of the loop kernel (traffic), of the dense Carry update, and of the multi-limb stage inside
# a 25FV47 solve (integer-pipe utilisation of the cooperative tier).
set -u
O=gpurun_out; mkdir -p $O
# gpurun brings back at most 64 MiB and a --set full report is 25 MB: export what is read here, drop the report
export_rep() { ncu -i $O/$1.ncu-rep --page raw --csv > $O/$1_raw.csv 2>/dev/null; ncu -i $O/$1.ncu-rep --page details --csv > $O/$1_details.csv 2>/dev/null; [ "${2:-}" = src ] && ncu -i $O/$1.ncu-rep --page source --csv > $O/$1_source.csv 2>/dev/null; rm -f $O/$1.ncu-rep; }
B="python bench.py --steps 2 --warmup 3 --cpu-pivots 10 --no-k1-dense"
timeout 300 $B > $O/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $B > $O/ncu_bench.log 2>&1
B1="python bench.py --steps 1 --warmup 3 --cpu-pivots 10 --no-k1-dense"
timeout 300 $B1 > $O/plain_bench1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_simplex_loop --launch-skip 7 --launch-count 1 -o $O/loop $B1 > $O/ncu_loop.log 2>&1
export_rep loop
timeout 300 python scripts/k1_dense.py 20000 3 > $O/k1_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_update$ -s 1 -c 1 -o $O/k1_dense python scripts/k1_dense.py 20000 3 > $O/k1_ncu.log 2>&1
export_rep k1_dense
timeout 300 python scripts/k1_dense.py 20000 8 | tail -1 > $O/k1_dense.json; cat $O/k1_dense.json
# multi-limb stage: the staged (non-persistent) driver launches k_update_big once per pivot; capture one launch deep in phase 1
S="python scripts/one_solve.py netlib_25FV47 4000 0"
timeout 600 $S > $O/big_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_update_big --launch-skip 3500 --launch-count 1 -o $O/update_big $S > $O/ncu_big.log 2>&1
export_rep update_big src
# the device arithmetic alone: instruction counts of ONE 128-bit and ONE 1600-bit x - a b (one thread), and of one done by a warp
for cfg in "128 0 1" "1600 0 1" "1600 3 1" "1600 0 32768"; do
  set -- $cfg
  timeout 120 python scripts/arith_time.py --one $1 $2 $3 > $O/arith_one_$1_$2_$3.log 2>&1 &&
  timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_uniform.sum \
      --clock-control none -k regex:k_debug_rat_op --csv --log-file $O/arith_ncu_$1_$2_$3.csv python scripts/arith_time.py --one $1 $2 $3 > /dev/null 2>&1
done
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_debug_rat_op -c 1 -o $O/arith_128 python scripts/arith_time.py --one 128 0 1 > /dev/null 2>&1
export_rep arith_128 src
ls -la $O
