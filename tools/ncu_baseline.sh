#!/bin/bash
# usage: tools/ncu_baseline.sh <tag>   (run under gpurun; writes gpurun_out/<tag>_*)
set -u
TAG=${1:-r01}
CMD="python bench.py --widths 1,8,17,32,40,63 --steps 2 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
QB="python tools/quick_bench.py 27 17,40"
$QB > gpurun_out/${TAG}_qb_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"decode_kernel|encode_kernel|analyze_kernel" -c 12 -o gpurun_out/${TAG}_prof $QB > gpurun_out/${TAG}_ncu2.log 2>&1
tail -3 gpurun_out/${TAG}_plain.log | cut -c1-600
tail -4 gpurun_out/${TAG}_qb_plain.log
ls -la gpurun_out/
