set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests7.log 2>&1
tail -8 gpurun_out/r1s9_gpu_tests7.log
for a in "filter 2048 2048 2048 u16" "filter 2048 2048 2048 u8" "filter 2048 2048 1024 f32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe7.log
done
timeout 300 python tools/brick_probe.py 2048 64 1 2>&1 | tail -2 | tee -a gpurun_out/r1s9_brick7.log
VKTB_BRICK_WORDS=0 timeout 300 python tools/brick_probe.py 2048 64 1 2>&1 | tail -2 | tee -a gpurun_out/r1s9_brick7.log
timeout 300 python tools/brick_probe.py 2048 64 0 2>&1 | tail -1 | tee -a gpurun_out/r1s9_brick7.log
VKTB_BRICK_WORDS=0 timeout 300 python tools/brick_probe.py 2048 64 0 2>&1 | tail -1 | tee -a gpurun_out/r1s9_brick7.log
timeout 300 python tools/brick_probe.py 2048 128 2 2>&1 | tail -1 | tee -a gpurun_out/r1s9_brick7.log
VKTB_BRICK_WORDS=0 timeout 300 python tools/brick_probe.py 2048 128 2 2>&1 | tail -1 | tee -a gpurun_out/r1s9_brick7.log
