set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:aggregatesKernel -c 1 -o gpurun_out/r1s9_agg_u16_v2 -f python tools/op_probe.py agg 2048 2048 1024 u16 > gpurun_out/ncu_agg2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:affineKernel -c 1 -o gpurun_out/r1s9_scale_u16_v2 -f python tools/op_probe.py scale 1024 1024 1024 u16 > gpurun_out/ncu_scale2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:affineKernel -c 1 -o gpurun_out/r1s9_rotate_u16_v2 -f python tools/op_probe.py rotate 1024 1024 1024 u16 > gpurun_out/ncu_rotate2.log 2>&1
ls -la gpurun_out | tail -5
