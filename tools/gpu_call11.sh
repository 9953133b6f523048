set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests11.log 2>&1
tail -8 gpurun_out/r1s9_gpu_tests11.log
for a in "rotate 2048 2048 2048 u16" "rotate 2048 2048 2048 u8" "rotate 2048 2048 1024 f32"; do
  timeout 300 python tools/op_probe.py $a 2>&1 | tail -1 | tee -a gpurun_out/r1s9_op_probe11.log
done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r1s9_bench11.json 2> gpurun_out/r1s9_bench11.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:Kernel -c 400 --csv \
  --log-file gpurun_out/r1s9_launches11.csv python bench.py --steps 5 --warmup 3 > gpurun_out/r1s9_ncu11.log 2>&1
tail -1 gpurun_out/r1s9_bench11.json
