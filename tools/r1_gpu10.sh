#!/bin/bash
mkdir -p gpurun_out
S=gpurun_out/status10.txt; : > $S
timeout 200 python bench.py --workload c4 --steps 500 --warmup 10 > gpurun_out/bench_c4_10.json 2> gpurun_out/bench_c4_10.err; echo "c4 rc=$?" >> $S
timeout 120 python bench.py --workload c2 --steps 200 > gpurun_out/bench_c2_10.json 2> gpurun_out/bench_c2_10.err; echo "c2 rc=$?" >> $S
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu10.log 2>&1; echo "gpu suite rc=$?" >> $S
cat $S; tail -2 gpurun_out/t_gpu10.log; tail -3 gpurun_out/bench_c4_10.err
