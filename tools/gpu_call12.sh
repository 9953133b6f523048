set -x
mkdir -p gpurun_out
timeout 300 python tools/op_probe.py sumrange 2048 2048 2048 u16 2>&1 | tail -1 | tee gpurun_out/r1s9_op_probe12.log
timeout 300 python tools/op_probe.py copyrange 2048 2048 2048 u16 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe12.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:arithRows -s 3 -c 1 -f -o gpurun_out/r1s9_sumrange_shifted \
  python tools/op_probe.py sumrange 2048 2048 1024 u16 > gpurun_out/r1s9_ncu12.log 2>&1
tail -3 gpurun_out/r1s9_ncu12.log
