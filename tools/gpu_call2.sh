set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests2.log 2>&1
tail -15 gpurun_out/r1s9_gpu_tests2.log
for a in "filter 2048 2048 2048 u16" "filter 2048 2048 1024 f32" "filter 2048 2048 2048 u8" "scale 2048 2048 2048 u16" "rotate 2048 2048 2048 u16" "scale 2048 2048 1024 f32" "rotate 2048 2048 1024 f32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe.log
done
# ncu: launch list + full capture of the aggregates and filter kernels (1024^3 keeps the replay short)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:aggregatesKernel -c 1 -o gpurun_out/r1s9_agg_u16 -f python tools/op_probe.py agg 2048 2048 1024 u16 > gpurun_out/ncu_agg.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:filter3MarchKernel -c 1 -o gpurun_out/r1s9_filter_u16 -f python tools/op_probe.py filter 1024 1024 1024 u16 > gpurun_out/ncu_filter.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:affineKernel -c 1 -o gpurun_out/r1s9_scale_u16 -f python tools/op_probe.py scale 1024 1024 1024 u16 > gpurun_out/ncu_scale.log 2>&1
ls -la gpurun_out | tail
