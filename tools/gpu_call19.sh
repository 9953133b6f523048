set -x
mkdir -p gpurun_out
timeout 300 python tools/op_probe.py sumrange 2048 2048 2048 u8 2>&1 | tail -1 | tee gpurun_out/r1s9_op_probe19.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:arithRowsPairTable -s 3 -c 1 -f -o gpurun_out/r1s9_sumrange_u8_pairrows2 \
  python tools/op_probe.py sumrange 2048 2048 1024 u8 > gpurun_out/r1s9_ncu19.log 2>&1
