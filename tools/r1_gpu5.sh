#!/bin/bash
mkdir -p gpurun_out
S=gpurun_out/status5.txt; : > $S
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu5.log 2>&1; echo "gpu suite rc=$?" >> $S
timeout 120 python tools/e2e_probe.py > gpurun_out/e2e_probe5.txt 2>&1; echo "probe rc=$?" >> $S
timeout 240 python bench.py --workload swarm --steps 400 --warmup 20 --no-cpu-baseline > gpurun_out/bench_swarm5.json 2> gpurun_out/bench_swarm5.err; echo "swarm rc=$?" >> $S
timeout 120 python bench.py --workload c2 --steps 200 > gpurun_out/bench_c2_5.json 2> gpurun_out/bench_c2_5.err; echo "c2 rc=$?" >> $S
cat $S; tail -3 gpurun_out/t_gpu5.log; cat gpurun_out/e2e_probe5.txt
