set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests20.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests20.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r1s9_smoke20.log
timeout 900 python bench.py > gpurun_out/r1s9_bench20.json 2> gpurun_out/r1s9_bench20.err
tail -c 600 gpurun_out/r1s9_bench20.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:arithRowsShifted -s 3 -c 1 -f -o gpurun_out/r1s9_sumrange_shifted_final \
  python tools/op_probe.py sumrange 2048 2048 1024 u16 > gpurun_out/r1s9_ncu20.log 2>&1
