#!/bin/bash
mkdir -p gpurun_out
S=gpurun_out/status2.txt; : > $S
timeout 300 python -m pytest tests/test_callers_gpu.py -x -q > gpurun_out/t_callers2.log 2>&1; echo "callers rc=$?" >> $S
timeout 240 python bench.py --workload swarm --steps 400 --warmup 20 > gpurun_out/bench_swarm2.json 2> gpurun_out/bench_swarm2.err; echo "swarm rc=$?" >> $S
timeout 240 python bench.py --workload swarm --steps 400 --warmup 20 --no-cpu-baseline > gpurun_out/bench_swarm2b.json 2> gpurun_out/bench_swarm2b.err; echo "swarm rc=$?" >> $S
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_swarm.csv python bench.py --workload swarm --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_swarm.log 2>&1; echo "ncu list rc=$?" >> $S
cat $S; tail -3 gpurun_out/t_callers2.log
