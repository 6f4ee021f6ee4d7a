#!/bin/bash
# One-box verification sequence (run under gpurun): new caller tests, swarm bench, whole GPU suite, headline bench.
mkdir -p gpurun_out
S=gpurun_out/status.txt; : > $S
timeout 300 python -m pytest tests/test_callers_gpu.py -x -q > gpurun_out/t_callers.log 2>&1; echo "callers rc=$?" >> $S
timeout 240 python bench.py --workload swarm --steps 400 --warmup 20 > gpurun_out/bench_swarm.json 2> gpurun_out/bench_swarm.err; echo "swarm rc=$?" >> $S
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?" >> $S
timeout 300 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?" >> $S
timeout 120 python bench.py --workload c2 --steps 200 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?" >> $S
timeout 120 python bench.py --workload swarm --robots 2 --rows 4000000 --steps 400 --warmup 20 > gpurun_out/bench_swarm_r2.json 2> gpurun_out/bench_swarm_r2.err; echo "swarm r2 rc=$?" >> $S
cat $S; tail -3 gpurun_out/t_callers.log; tail -3 gpurun_out/t_gpu.log
