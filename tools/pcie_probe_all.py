"""Host<->device ceiling of the whole box: tools/pcie_probe.py's three measurements (H2D, D2H, both directions at once,
pinned 1 GiB buffers) on ALL visible GPUs at the same time, one process per GPU started on a common wall-clock tick.
The aggregate `both` figure is what bench.py's end-to-end (host buffer) leg can reach at N GPUs: its step is H2D of the
raw columns + D2H of payloads / decoded vectors, both directions busy.
usage: python tools/pcie_probe_all.py [n_gpus] -> one JSON line"""
import json
import os
import subprocess
import sys
import time

CHILD = r'''
import json, os, sys, time, torch
start = float(sys.argv[1])
n = 1 << 30
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def h2d():
    with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
def both():
    h2d(); d2h()
def timed(fn, tick, reps=8):
    fn(); torch.cuda.synchronize()
    while time.time() < tick: pass
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
out = {}
for k, (name, fn, vol) in enumerate((("h2d", h2d, n), ("d2h", d2h, n), ("both", both, 2 * n))):
    out[name] = vol / timed(fn, start + 4.0 * k) / 1e9
print(json.dumps(out))
'''


def main():
    import torch
    n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
    start = time.time() + 25.0  # torch import + 2 GiB of pinned memory per process
    procs = []
    for g in range(n):
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(g))
        procs.append(subprocess.Popen([sys.executable, "-c", CHILD, str(start)], env=env, stdout=subprocess.PIPE, text=True))
    per = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
    agg = {k: round(sum(r[k] for r in per), 1) for k in ("h2d", "d2h", "both")}
    print(json.dumps({"gpus": n, "aggregate_GBs": agg, "per_gpu_GBs": [{k: round(v, 1) for k, v in r.items()} for r in per],
                      "note": "pinned 1 GiB copies, all GPUs concurrently; `both` = H2D and D2H streams busy at once (sum of both directions)"}))


main()
