set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests.log
for a in "2048 2048 2048 u16" "2048 2048 2048 u16 -1000 3000" "2048 2048 2048 u8" "2048 2048 1024 f32" "2048 2048 1024 u32" "2048 2048 2048 u8 -1000 3000"; do
  timeout 300 python tools/agg_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_agg_probe.log
done
