set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests5.log 2>&1
tail -15 gpurun_out/r1s9_gpu_tests5.log
for a in "agg 2048 2048 2048 u16" "agg 2048 2048 2048 u8" "agg 2048 2048 1024 f32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe5.log
done
timeout 900 python bench.py --steps 5 --warmup 3 --no-e2e > gpurun_out/r1s9_bench5.json 2> gpurun_out/r1s9_bench5.err
tail -c 600 gpurun_out/r1s9_bench5.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1s9_bench5.json'))
for k,v in d.get('ops',{}).items(): print(f"{v['ms']:9.3f} ms {100*v['frac_of_measured_peak']:6.1f}%  {k}")
PY
