set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests10.log 2>&1
tail -8 gpurun_out/r1s9_gpu_tests10.log
for a in "scale 2048 2048 2048 u16" "scale 2048 2048 2048 u8" "scale 2048 2048 1024 u32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe10.log
  VKTB_SCALE_ROWS=0 timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe10.log
done
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
