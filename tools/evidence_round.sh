#!/bin/bash
# usage (under gpurun, one GPU): tools/evidence_round.sh <tag> [sections]
# Per-round evidence that is not in the bench line (VERDICT r1 "next" item 9): GPU test log, per-kernel developer
# loops for the narrow physical types, the single-column filter, Q6/Q1 and fetch_row, ncu launch lists of the same
# commands, and compute-sanitizer memcheck / racecheck over the parity and edge tests.  Everything lands in
# gpurun_out/<tag>_*; numbers printed under ncu or the sanitizer are never bench values.
set -u
TAG=${1:-r02}
SECTIONS=${2:-tests,loops,ncu,sanitize}
OUT=gpurun_out
mkdir -p $OUT
has() { [[ ",$SECTIONS," == *",$1,"* ]]; }

if has tests; then
  python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest.log
fi
if has loops; then
  for t in uint8 int16 uint32; do
    python tools/quick_bench.py 28 3,7,13 $t > $OUT/${TAG}_qb_$t.log 2>&1 || tail -3 $OUT/${TAG}_qb_$t.log
  done
  python tools/quick_bench.py 27 1,4,8,12,17,24,32,40,48,63 > $OUT/${TAG}_qb_int64.log 2>&1
  python tools/filter_bench.py 27 4,8,12,17,24,32,40,63 > $OUT/${TAG}_filter_bench.log 2>&1 || tail -3 $OUT/${TAG}_filter_bench.log
  python tools/query_bench.py 10 nocpu > $OUT/${TAG}_query_bench.log 2>&1 || tail -3 $OUT/${TAG}_query_bench.log
  python tools/fetch_bench.py > $OUT/${TAG}_fetch_bench.log 2>&1 || tail -3 $OUT/${TAG}_fetch_bench.log
  tail -n 3 $OUT/${TAG}_qb_uint8.log $OUT/${TAG}_filter_bench.log $OUT/${TAG}_query_bench.log $OUT/${TAG}_fetch_bench.log | cut -c1-400
fi
if has ncu; then
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum \
      --clock-control none -k regex:"fetch_rows_kernel" -c 15 --csv --log-file $OUT/${TAG}_fetch_ncu.csv \
      python tools/fetch_bench.py > $OUT/${TAG}_fetch_ncu.log 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_qb_uint8_launches.csv \
      python tools/quick_bench.py 28 7 uint8 > /dev/null 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_qb_uint32_launches.csv \
      python tools/quick_bench.py 28 13 uint32 > /dev/null 2>&1
  ncu --set full --clock-control none -k regex:"decode_kernel|pack_staged_kernel" -c 6 -o $OUT/${TAG}_narrow_prof \
      python tools/quick_bench.py 28 7 uint8 > $OUT/${TAG}_narrow_ncu.log 2>&1
fi
if has sanitize; then
  timeout 600 compute-sanitizer --tool memcheck --print-limit 20 --log-file $OUT/${TAG}_memcheck.log \
      python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_gpu_select.py -m gpu -x -q \
      > $OUT/${TAG}_memcheck_pytest.log 2>&1; echo "memcheck rc=$?"
  tail -3 $OUT/${TAG}_memcheck.log; tail -2 $OUT/${TAG}_memcheck_pytest.log
  timeout 420 compute-sanitizer --tool racecheck --print-limit 20 --log-file $OUT/${TAG}_racecheck.log \
      python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "int64 or uint8 or INT64 or UINT8 or roundtrip or filter" \
      > $OUT/${TAG}_racecheck_pytest.log 2>&1; echo "racecheck rc=$?"
  tail -3 $OUT/${TAG}_racecheck.log; tail -2 $OUT/${TAG}_racecheck_pytest.log
fi
ls -la $OUT | grep ${TAG}
