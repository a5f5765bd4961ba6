#!/bin/bash
# The round's ncu passes (one GPU; each capture only after the same command exited 0 without ncu): launch list of the
# bench command, --set full of the loop kernel (traffic), of the dense Carry update, and of the multi-limb stage inside
# a 25FV47 solve (integer-pipe utilisation of the cooperative tier).
set -u
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 2 --warmup 3 --cpu-pivots 10 --no-k1-dense"
timeout 300 $B > $O/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $B > $O/ncu_bench.log 2>&1
B1="python bench.py --steps 1 --warmup 3 --cpu-pivots 10 --no-k1-dense"
timeout 300 $B1 > $O/plain_bench1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_simplex_loop --launch-skip 7 --launch-count 1 -o $O/loop $B1 > $O/ncu_loop.log 2>&1
timeout 300 python scripts/k1_dense.py 20000 3 > $O/k1_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_update$ -s 1 -c 1 -o $O/k1_dense python scripts/k1_dense.py 20000 3 > $O/k1_ncu.log 2>&1
timeout 300 python scripts/k1_dense.py 20000 8 | tail -1 > $O/k1_dense.json; cat $O/k1_dense.json
# multi-limb stage: the staged (non-persistent) driver launches k_update_big once per pivot; capture one launch deep in phase 1
S="python scripts/one_solve.py netlib_25FV47 4000 0"
timeout 600 $S > $O/big_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_update_big --launch-skip 3500 --launch-count 1 -o $O/update_big $S > $O/ncu_big.log 2>&1
ls -la $O
