set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q -k "range or arith or shifted or walking") > gpurun_out/r1s9_gpu_tests25.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests25.log
for e in 0 1; do
for a in "sumrange 2048 2048 2048 u16" "sumrange 2048 2048 1024 f32" "sumrange 2048 2048 1024 u32"; do
  VKTB_ARITH_EXCHANGE=$e timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe25.log
done
done
