#!/bin/bash
# One-box verification sequence (run under gpurun): whole GPU suite, smoke(), headline bench, swarm bench.
mkdir -p gpurun_out
S=gpurun_out/status.txt; : > $S
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?" >> $S
timeout 60 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> $S
timeout 300 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?" >> $S
timeout 200 python bench.py --workload swarm --steps 400 --warmup 20 > gpurun_out/bench_swarm.json 2> gpurun_out/bench_swarm.err; echo "swarm rc=$?" >> $S
cat $S; tail -2 gpurun_out/t_gpu.log; tail -1 gpurun_out/smoke.log
