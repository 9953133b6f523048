set -x
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r1s9_gpu_tests28.log 2>&1
tail -5 gpurun_out/r1s9_gpu_tests28.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r1s9_smoke28.log
timeout 900 python bench.py > gpurun_out/r1s9_bench28.json 2> gpurun_out/r1s9_bench28.err
tail -c 300 gpurun_out/r1s9_bench28.json
