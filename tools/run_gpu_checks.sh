#!/bin/bash
# The one GPU-side runner.  Everything it writes goes to gpurun_out/ (scratch; summaries worth keeping are copied to
# profiles/ by hand with tools/ncu_summary.py).  Run on a B200 box through gpurun:
#
#   gpurun --timeout 900 -- 'bash tools/run_gpu_checks.sh TAG [steps...]'
#
# steps (default: tests smoke bench):
#   tests      python -m pytest tests -m gpu -x -q
#   smoke      __graft_entry__.smoke()
#   bench      python bench.py                       (N = 1, the driver's command)
#   refbench   python bench.py --impl reference
#   launches   ncu launch list of the bench command without its host-streaming leg (only after `bench` exited 0 without ncu)
#   ncu:REGEX  one `ncu --set full` capture of the kernels matching REGEX in `bench.py --ops 1 --steps 1 --warmup 1`
#   probe:ARGS python tools/op_probe.py ARGS          (ARGS with commas for spaces, e.g. probe:sumrange,2040,2040,2040,u8)
set -u
TAG=${1:-run}; shift || true
STEPS=${*:-tests smoke bench}
OUT=gpurun_out
mkdir -p $OUT
for s in $STEPS; do
  case $s in
    tests)    timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?" ;;
    smoke)    timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" ;;
    bench)    timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?" ;;
    refbench) timeout 600 python bench.py --impl reference > $OUT/${TAG}_refbench.json 2> $OUT/${TAG}_refbench.err; echo "refbench rc=$?" ;;
    launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:Kernel|vktb_' -c 1000 --csv \
                --log-file $OUT/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e > $OUT/${TAG}_launches.log 2>&1; echo "launches rc=$?" ;;
    ncu:*)    timeout 900 ncu --set full --clock-control none --import-source on -k regex:${s#ncu:} -c 3 -f \
                -o $OUT/${TAG}_$(echo ${s#ncu:} | tr -c 'A-Za-z0-9' _) python bench.py --steps 1 --warmup 1 --no-e2e \
                > $OUT/${TAG}_ncu.log 2>&1; echo "ncu rc=$?" ;;
    probe:*)  timeout 300 python tools/op_probe.py $(echo ${s#probe:} | tr ',' ' ') >> $OUT/${TAG}_probe.log 2>&1; echo "probe rc=$?" ;;
    *)        echo "unknown step $s" ;;
  esac
done
tail -n 5 $OUT/${TAG}_*.log 2>/dev/null | tail -n 60
