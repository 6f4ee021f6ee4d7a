#!/bin/bash
mkdir -p gpurun_out
S=gpurun_out/status_final2.txt; : > $S
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/t_gpu_final2.log 2>&1; echo "gpu suite rc=$?" >> $S
timeout 60 python __graft_entry__.py smoke > gpurun_out/smoke_final2.log 2>&1; echo "smoke rc=$?" >> $S
timeout 300 python bench.py > gpurun_out/bench_c3_final2.json 2> gpurun_out/bench_c3_final2.err; echo "c3 rc=$?" >> $S
timeout 200 python bench.py --workload swarm --steps 400 --warmup 20 > gpurun_out/bench_swarm_final2.json 2> gpurun_out/bench_swarm_final2.err; echo "swarm rc=$?" >> $S
cat $S; tail -2 gpurun_out/t_gpu_final2.log; tail -1 gpurun_out/smoke_final2.log
