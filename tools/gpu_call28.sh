set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests28.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests28.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r1s9_smoke28.log
timeout 900 python bench.py > gpurun_out/r1s9_bench28.json 2> gpurun_out/r1s9_bench28.err
tail -c 300 gpurun_out/r1s9_bench28.json
for a in "sumrange 2048 2048 2048 u16" "sumrange 2048 2048 1024 f32" "sumrange 2048 2048 1024 u32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe28.log
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:arithRowsShifted -s 3 -c 1 -f -o gpurun_out/r1s9_sumrange_shifted_final2 \
  python tools/op_probe.py sumrange 2048 2048 1024 u16 > gpurun_out/r1s9_ncu28.log 2>&1
