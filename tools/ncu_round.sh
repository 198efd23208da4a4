#!/bin/bash
# usage (under gpurun, one GPU): tools/ncu_round.sh <tag>
# Writes gpurun_out/<tag>_*: plain bench run, ncu launch list of the same command, dram traffic of the dominant
# kernel over one whole step, and one `--set full` capture of the hot kernels.  Numbers printed under ncu are never
# bench values.
set -u
TAG=${1:-r01}
OUT=gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --configs ''"
$BENCH > $OUT/${TAG}_plain.log 2>&1 || { echo "plain bench failed"; tail -5 $OUT/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT/${TAG}_launches.csv \
    $BENCH > $OUT/${TAG}_ncu_launches.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"decode_kernel" -c 400 --csv --log-file $OUT/${TAG}_traffic.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --configs '' > $OUT/${TAG}_ncu_traffic.log 2>&1
QB="python tools/quick_bench.py 27 17,40"
$QB > $OUT/${TAG}_qb_plain.log 2>&1 &&
ncu --set full --clock-control none \
    -k regex:"decode_kernel|pack_staged_kernel|analyze_kernel|scan_sum_fields|layout_kernel" -c 20 -o $OUT/${TAG}_prof \
    $QB > $OUT/${TAG}_ncu_full.log 2>&1
tail -c 400 $OUT/${TAG}_plain.log; echo
tail -2 $OUT/${TAG}_qb_plain.log | cut -c1-300
ls -la $OUT | grep ${TAG}
