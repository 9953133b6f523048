set -x
mkdir -p gpurun_out
for e in 0 1; do
  for a in "sumrange8 2048 2048 2048 u16" "sumrange8 2048 2048 1024 f32"; do
    VKTB_ARITH_WALK_ALIGNED=$e timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe22.log
  done
done
