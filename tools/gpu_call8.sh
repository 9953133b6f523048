set -x
mkdir -p gpurun_out
nvidia-smi -L
(time timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q) > gpurun_out/r1s9b_gpu_tests_n2.log 2>&1
tail -6 gpurun_out/r1s9b_gpu_tests_n2.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --shard-edge 2048 > gpurun_out/r1s9b_bench_n2.json 2> gpurun_out/r1s9b_bench_n2.err
tail -c 400 gpurun_out/r1s9b_bench_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1s9b_bench_n2.json'))
print(d['metric'], d['value'], d['n_gpus'], d['ms_per_step'])
for k,v in d.get('sharded',{}).items():
    print(k, {a:b for a,b in v.items() if a in ('ms','gvoxels_s','frac_of_measured_peak')})
PY
