B="python bench.py --no-cpu-baseline --no-torch-bar --no-e2e --steps 5"
for cfg in "0 10" "2 10" "18 10" "19 10" "18 32"; do
  set -- $cfg
  echo "== dbg=$1 k=$2" >> gpurun_out/screen_probe2.txt
  CSLAM_NNS_K2_DBG=$1 timeout 200 $B --k $2 2>> gpurun_out/screen_probe2.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms_per_step %.3f kernel_ms %.3f value %.0f' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['value']))" >> gpurun_out/screen_probe2.txt 2>&1
done
cat gpurun_out/screen_probe2.txt
