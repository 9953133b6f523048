set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests13.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests13.log
for e in 1 0; do
  VKTB_ARITH_EXCHANGE=$e timeout 300 python tools/op_probe.py sumrange 2048 2048 2048 u16 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe13.log
  VKTB_ARITH_EXCHANGE=$e timeout 300 python tools/op_probe.py sumrange 2048 2048 2048 u8 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe13.log
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:arithRows -s 3 -c 1 -f -o gpurun_out/r1s9_sumrange_exchange \
  python tools/op_probe.py sumrange 2048 2048 1024 u16 > gpurun_out/r1s9_ncu13.log 2>&1
tail -3 gpurun_out/r1s9_ncu13.log
