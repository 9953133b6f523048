set -x
mkdir -p gpurun_out
timeout 1500 python bench.py > gpurun_out/r1s9_final_bench_n1.json 2> gpurun_out/r1s9_final_bench_n1.err
tail -c 300 gpurun_out/r1s9_final_bench_n1.err
python -c "
import json
d=json.load(open('gpurun_out/r1s9_final_bench_n1.json'))
print({k:d[k] for k in ('metric','value','ms_per_step','gpu_launches')}, d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline'])
"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1s9_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --ops 1 > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/ncu_launches.log | cut -c1-300
