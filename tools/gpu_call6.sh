set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q -k "affine or golden or host or filter") > gpurun_out/r1s9_gpu_tests6.log 2>&1
tail -8 gpurun_out/r1s9_gpu_tests6.log
for a in "filter 2048 2048 2048 u16" "filter 2048 2048 2048 u8" "filter 2048 2048 1024 f32" "rotate 2048 2048 2048 u16" "rotate 2048 2048 1024 f32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe6.log
done
VKTB_FILTER_PAIRS=0 timeout 300 python tools/op_probe.py filter 2048 2048 2048 u16 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe6.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:filter3PairKernel -c 1 -o gpurun_out/r1s9_filter_pair_u16 -f python tools/op_probe.py filter 1024 1024 1024 u16 > gpurun_out/ncu_filter2.log 2>&1
