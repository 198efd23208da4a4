"""Mint TPC-H golden inputs + REFERENCE answers with the L2 driver (the unmodified reference, oracle/l2).

Runs where /root/reference exists (needs oracle/l2/build_drivers.sh to have built oracle/_ref/l2/l2_driver_ref):
  * SF0.01 -> tests/golden/tpch_sf0_01.npz (lineitem's integer columns as dbgen produced them, + the l_rfls group code)
              tests/golden/tpch_sf0_01.json (Q1 / Q6 answers, sizes before/after CompactAllSegments) -- committed;
  * SF0.1, SF1 -> oracle/_ref/l2/tpch_sf{0.1,1}/ (flat column files + answers.json; git-ignored, travels to the GPU box).
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
L2 = os.path.join(ROOT, "oracle", "_ref", "l2")
COLUMNS = {"l_orderkey": "i32", "l_partkey": "i32", "l_suppkey": "i32", "l_linenumber": "i32", "l_quantity": "i32",
           "l_extendedprice": "i64", "l_discount": "i64", "l_tax": "i64", "l_shipdate": "i32", "l_commitdate": "i32",
           "l_receiptdate": "i32", "l_rfls": "u8"}
DTYPES = {"i32": np.int32, "i64": np.int64, "u8": np.uint8}


def run(sf: str) -> tuple[dict, str]:
    out = os.path.join(L2, f"tpch_sf{sf}")
    os.makedirs(out, exist_ok=True)
    line = subprocess.check_output([os.path.join(L2, "l2_driver_ref"), "tpch", sf, out, "1", "3"]).decode().splitlines()[-1]
    ans = json.loads(line)
    json.dump(ans, open(os.path.join(out, "answers.json"), "w"), indent=1)
    return ans, out


def load_columns(directory: str) -> dict:
    return {name: np.fromfile(os.path.join(directory, f"{name}.{ext}"), dtype=DTYPES[ext]) for name, ext in COLUMNS.items()}


if __name__ == "__main__":
    sfs = sys.argv[1:] or ["0.01", "0.1", "1"]
    for sf in sfs:
        ans, out = run(sf)
        print(sf, ans["rows"], ans["q6_revenue"])
        if sf == "0.01":
            cols = load_columns(out)
            np.savez_compressed(os.path.join(ROOT, "tests", "golden", "tpch_sf0_01.npz"), **cols)
            json.dump(ans, open(os.path.join(ROOT, "tests", "golden", "tpch_sf0_01.json"), "w"), indent=1)
