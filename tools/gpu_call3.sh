set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests3.log 2>&1
tail -15 gpurun_out/r1s9_gpu_tests3.log
for a in "filter 2048 2048 2048 u16" "filter 2048 2048 2048 u8" "scale 2048 2048 2048 u16" "rotate 2048 2048 2048 u16" "scale 2048 2048 1024 f32" "rotate 2048 2048 1024 f32" "scale 2048 2048 2048 u8" "agg 2048 2048 2048 u16" "agg 2048 2048 2048 u8"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe3.log
done
df -h /dev/shm | tail -1
timeout 600 python tools/io_probe.py 8 /dev/shm 2>&1 | tail -2 | tee -a gpurun_out/r1s9_io_probe.log
timeout 600 python tools/io_probe.py 8 /tmp 2>&1 | tail -2 | tee -a gpurun_out/r1s9_io_probe.log
