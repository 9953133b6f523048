set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests24.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests24.log
for a in "copyrange 2048 2048 2048 u16" "copyrange 2048 2048 2048 u8" "copyrange 2048 2048 1024 f32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe24.log
done
