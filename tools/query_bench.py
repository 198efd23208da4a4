"""Developer loop for the fused multi-column scan: TPC-H SF10 Q6 / Q1 device times (tools/workloads.config4_tpch)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckdb_adaptive_compression_b200 as adacom  # noqa: E402
from tools import workloads  # noqa: E402

ctx = adacom.Context(0)
peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
sf = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
r = workloads.config4_tpch(ctx, peak, os.cpu_count() or 1, sf=sf, cpu_row_groups=0 if len(sys.argv) > 2 else 24)
print(json.dumps({"q6": r["q6"], "q1": r["q1"], "cpu": r.get("cpu_baseline", {}).get("value")}))
