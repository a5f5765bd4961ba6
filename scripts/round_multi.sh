#!/bin/bash
# One `gpurun --gpus N` call: the multi-GPU parity tests and the bench lines at N GPUs (one process per GPU, torchrun).
set -u
N=${1:-2}; O=gpurun_out; mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(time timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -m gpu) > $O/gpu_multi_tests_g$N.log 2>&1; tail -3 $O/gpu_multi_tests_g$N.log
timeout 600 $T --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_g$N.json 2> $O/bench_g$N.err; cut -c1-300 $O/bench_g$N.json; tail -2 $O/bench_g$N.err
timeout 600 $T --master-port 29512 bench.py --gpus $N --workload miplib_50v-10 --steps 3 --warmup 3 > $O/bench_g${N}_50v-10.json 2> $O/bench_g${N}_50v-10.err; cut -c1-300 $O/bench_g${N}_50v-10.json; tail -2 $O/bench_g${N}_50v-10.err
timeout 300 $T --master-port 29513 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > $O/bench_reference_g$N.json 2> $O/bench_reference_g$N.err; cut -c1-200 $O/bench_reference_g$N.json
