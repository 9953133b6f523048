set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:arithRowsPairTable -s 3 -c 1 -f -o gpurun_out/r1s9_sumrange_u8_pairrows \
  python tools/op_probe.py sumrange 2048 2048 1024 u8 > gpurun_out/r1s9_ncu18.log 2>&1
tail -3 gpurun_out/r1s9_ncu18.log
