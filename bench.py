#!/usr/bin/env python
"""bench.py -- succinct codec throughput on B200 (BASELINE.json configs[1]).

Workload (per GPU): INT64 columns of 2^27 values (1 GiB raw), value = splitmix64(seed + global_index) &
(2^b - 1) for every bit width b = 1..64, DuckDB segmentation (row groups of 122 880 rows -> segments of
32767/32767/32767/24579 values).  One STEP = the whole width sweep, and for every width the five hot-path
operations of SURVEY.md section 8(a):

    analyze   (SuccinctAppendLoop min/max)            algorithmic HBM bytes  n*S
    encode    (BitCompressFromSuccinct)                                      n*S + P
    scan      (SuccinctScan, all vectors -> flat)                            P + n*S
    scan+SUM  (fused decode + 128-bit SUM)                                   P
    uncompact (UncompressSuccinct)                                           P + n*S

`value` = logical bytes through the codec per second = ops * n * S / t (the reference's "n*sizeof(T)/t"),
whole job over all ranks, inputs resident in HBM.  `e2e` = the same five operations through the C ABI with
HOST buffers (pinned): raw values H2D, packed payload + metadata D2H, decoded vectors D2H, aggregate D2H.
`roofline` = algorithmic bytes / CUDA-event time of the dominant kernel (decode_kernel, flat mode) against
MEASURED_PEAKS.json.  `cpu_baseline` / `--impl reference` = the reference's own SDSL code path
(oracle/_ref) timed on the box's host cores on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

OPS = ("analyze", "encode", "scan", "scan_sum", "uncompact")
METRIC = "succinct decode+scan GB/s and encode GB/s vs HBM roofline at 1/2/4/8 B200"
WORKLOAD = "uniform random int64 columns, bit-width sweep 1-64, 1 GiB/GPU, encode + full scan + SUM"
SEED0 = 0xADAC0000


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2-values", type=int, default=27, help="values per GPU per width (2^27 = 1 GiB of INT64)")
    ap.add_argument("--widths", default="1-64")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-log2-values", type=int, default=21)
    ap.add_argument("--detail", default="", help="write the per-width table to this JSON file")
    ap.add_argument("--configs", default="1,3,4,5,codecs,vtable", help="the other BASELINE.json workload shapes to run after the "
                    "headline sweep (1 Zipf point filters, 3 skew sweep + policy epochs, 4 TPC-H SF10 Q6/Q1, "
                    "5 TPC-H SF100 sharded Q6 over NCCL, codecs = DuckDB bitpacking / RLE, vtable = the drop-in "
                    "CompressionFunction slots driven like ColumnData from C++); '' = none")
    return ap.parse_args()


def parse_widths(spec: str):
    out = []
    for part in spec.split(","):
        if "-" in part:
            a, b = part.split("-")
            out += list(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    return out


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's own SDSL code path (oracle/_ref) or the C port, all host threads
# ---------------------------------------------------------------------------------------------------

def cpu_sweep(widths, log2_values, threads, steps=1, warmup=0):
    """One CPU 'step' = the same five operations per width on a bounded sample of 2^log2_values values."""
    import oracle
    from tests.cases import splitmix64
    kind = "reference" if oracle.have_reference() else "port"
    o = oracle.Oracle(kind)
    n = 1 << log2_values
    base = splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(SEED0))
    per_step = []
    for it in range(warmup + steps):
        t_ops = dict.fromkeys(OPS, 0.0)
        for b in widths:
            mask = np.uint64((1 << b) - 1) if b < 64 else np.uint64(0xFFFFFFFFFFFFFFFF)
            vals = (base & mask).view(np.int64)
            r = o.bench_column(vals, 32767, threads)
            t_ops["analyze"] += r["analyze_s"]
            t_ops["encode"] += r["encode_s"]
            t_ops["scan"] += r["scan_s"]
            t_ops["scan_sum"] += r["sum_s"]
            t_ops["uncompact"] += r["uncompact_s"]
        if it >= warmup:
            per_step.append(t_ops)
    total_s = sum(sum(t.values()) for t in per_step) / len(per_step)
    logical = len(OPS) * n * 8 * len(widths)
    ops = {k: round(n * 8 * len(widths) / (sum(t[k] for t in per_step) / len(per_step)) / 1e9, 3) for k in OPS}
    return dict(value=logical / total_s / 1e9, seconds_per_step=total_s, kind=kind, cores=threads,
                sample=f"{len(widths)} widths x 2^{log2_values} INT64 values per step, segments of 32767, "
                       f"one segment per task dealt round-robin over {threads} threads, 5 ops per width",
                ops_GBs=ops)


def shared_config(args, n_segments=None):
    """the `config` object, identical in both arms (--impl ours / reference): names the workload and states how the CPU
    arm samples it"""
    cfg = {"workload": WORKLOAD, "widths": args.widths, "values_per_gpu_per_width": 1 << args.log2_values,
           "ops_per_width": list(OPS),
           "cpu_arm_sample": f"2^{args.cpu_log2_values} values per width per step (same generator, same segmentation, same "
                             f"five ops; the GPU arm runs 2^{args.log2_values})",
           "value_definition": "logical bytes = ops * n * 8 per width, / time",
           "l2": "every op streams >= 1 GiB (> 126 MB L2) between reuses of any line"}
    return cfg


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    widths = parse_widths(args.widths)
    threads = os.cpu_count() or 1
    t0 = time.time()
    # bounded: about 1 s of CPU work per step on a 16-core host at 2^21 values per width
    res = cpu_sweep(widths, args.cpu_log2_values, threads, steps=max(args.steps, 1), warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": round(res["value"], 3), "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(res["seconds_per_step"] * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": shared_config(args),
        "cpu_baseline": {"value": round(res["value"], 3), "unit": "GB/s", "cores": threads, "kind": res["kind"],
                         "sample": res["sample"], "ops_GBs": res["ops_GBs"]},
        "e2e": {"value": round(res["value"], 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": round(time.time() - t0, 1),
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------

def host_mem_available_bytes() -> int:
    avail = 0
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    avail = int(ln.split()[1]) * 1024
    except Exception:
        pass
    for path in ("/sys/fs/cgroup/memory.max", "/sys/fs/cgroup/memory/memory.limit_in_bytes"):
        try:
            v = open(path).read().strip()
            if v.isdigit():
                lim = int(v)
                used = 0
                for up in ("/sys/fs/cgroup/memory.current", "/sys/fs/cgroup/memory/memory.usage_in_bytes"):
                    try:
                        used = int(open(up).read().strip())
                        break
                    except Exception:
                        pass
                avail = min(avail, lim - used) if avail else lim - used
        except Exception:
            pass
    return avail


def run_ours(args):
    import torch
    import torch.distributed as dist
    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200 import sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    widths = parse_widths(args.widths)
    n = 1 << args.log2_values
    S = 8
    dtype = np.dtype(np.int64)
    counts = adacom.row_group_segment_counts(n, dtype)
    ctx = adacom.Context(local_rank)

    # ---- resident inputs: one raw column per width, generated on the device ----
    cols = {}
    for b in widths:
        col = ctx.column(dtype, counts)
        col.fill_splitmix(SEED0 + b, rank * n, (1 << b) - 1)
        cols[b] = col
    flat = torch.empty(n * S + 256, dtype=torch.uint8, device=dev)
    aggs = torch.zeros(len(widths) * 4, dtype=torch.int64, device=dev)
    ctx.synchronize()

    n_ev = len(OPS) + 1
    # NCCL work is enqueued on the codec's own stream (an ExternalStream view of it): the all-gather runs right behind
    # the last kernel of the sweep, without a host synchronisation in between
    codec_stream = torch.cuda.ExternalStream(ctx.stream_ptr, device=dev)

    def step(events=None):
        for wi, b in enumerate(widths):
            col = cols[b]
            ev = events[wi] if events is not None else None
            if ev: ctx.record(ev[0])
            col.analyze()
            if ev: ctx.record(ev[1])
            col.encode()
            if ev: ctx.record(ev[2])
            col.decode(device_out=flat.data_ptr())
            if ev: ctx.record(ev[3])
            col.scan_sum_async(aggs.data_ptr() + wi * 32)
            if ev: ctx.record(ev[4])
            col.uncompact()
            if ev: ctx.record(ev[5])
        if world > 1:
            # the only cross-GPU exchange of the path: 64 x {sum_lo, sum_hi, count} per rank over NCCL
            with torch.cuda.stream(codec_stream):
                return sharding.all_gather_aggregates(aggs)
        return None

    def barrier():
        ctx.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()

    events = [[[ctx.event() for _ in range(n_ev)] for _ in widths] for _ in range(args.steps)]
    e_start, e_stop = ctx.event(), ctx.event()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count
    barrier()
    ctx.record(e_start)
    gathered = None
    for k in range(args.steps):
        gathered = step(events[k])
    ctx.record(e_stop)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launch_count - launches0
    total_ms = ctx.elapsed_ms(e_start, e_stop)
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt)
        launches = int(lt.item())
    ms_per_step = total_ms / args.steps

    # ---- per-op / per-width device times from the in-region events ----
    packed_bytes = {}
    for b in widths:  # widths are a pure function of b here; take P from one more encode outside the timing
        cols[b].compact()
        packed_bytes[b] = sum(m.words for m in cols[b].meta()) * 8
    op_ms = {op: np.zeros(len(widths)) for op in OPS}
    for k in range(args.steps):
        for wi in range(len(widths)):
            ev = events[k][wi]
            for oi, op in enumerate(OPS):
                op_ms[op][wi] += ctx.elapsed_ms(ev[oi], ev[oi + 1]) / args.steps
    raw_bytes = n * S
    alg = {"analyze": lambda P: raw_bytes, "encode": lambda P: raw_bytes + P, "scan": lambda P: P + raw_bytes,
           "scan_sum": lambda P: P, "uncompact": lambda P: P + raw_bytes}
    # b = 64 is not re-packed by the reference (Q1 domain): its encode moves no bytes, scan/uncompact copy raw
    def alg_bytes(op, b):
        P = packed_bytes[b]
        if P == 0:
            return {"analyze": raw_bytes, "encode": 0, "scan": 2 * raw_bytes, "scan_sum": raw_bytes,
                    "uncompact": 0}[op]
        return alg[op](P)
    peak, peak_src = measured_peak()
    detail = {"widths": widths, "packed_bytes": [packed_bytes[b] for b in widths], "ops": {}}
    op_summary = {}
    for op in OPS:
        tot_bytes = sum(alg_bytes(op, b) for b in widths)
        tot_ms = float(op_ms[op].sum())
        op_summary[op] = {"ms_per_step": round(tot_ms, 4),
                          "logical_GBs": round(raw_bytes * len(widths) / (tot_ms * 1e-3) / 1e9, 1),
                          "algorithmic_GBs": round(tot_bytes / (tot_ms * 1e-3) / 1e9, 1),
                          "frac_of_peak": round(tot_bytes / (tot_ms * 1e-3) / 1e9 / peak, 4)}
        detail["ops"][op] = {"ms": [round(float(x), 5) for x in op_ms[op]],
                             "algorithmic_GBs": [round(alg_bytes(op, b) / (op_ms[op][wi] * 1e-3) / 1e9, 1)
                                                 for wi, b in enumerate(widths)]}
    # dominant kernel: decode_kernel<8, false, flat> runs inside `scan` and `uncompact` (one launch each)
    dom_ms = float(op_ms["scan"].sum())
    dom_bytes = sum(alg_bytes("scan", b) for b in widths)
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of that kernel per launch, from the committed ncu capture of this
    # command (tools/ncu_round.sh -> tools/traffic_from_ncu.py); null when no capture covers the full sweep
    traffic = None
    try:
        if args.widths == "1-64" and args.log2_values == 27:
            with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
                traffic = int(json.load(f)["traffic_bytes_per_launch"])
    except Exception:
        traffic = None
    roofline = {"bound": "hbm", "kernel": "decode_kernel<8,false,flat> (scan -> flat vectors)",
                "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                "traffic": traffic, "peak_source": peak_src,
                "launches_per_step": len(widths), "avg_launch_ms": round(dom_ms / len(widths), 5),
                "algorithmic_bytes_per_launch": int(dom_bytes / len(widths))}

    # ---- sanity: fused SUM == sum of the decoded vectors == sum of the generator (one width, exact) ----
    check_b = widths[len(widths) // 2]
    cols[check_b].decode(device_out=flat.data_ptr())
    ctx.synchronize()
    dec = flat[: n * S].view(torch.int64)
    got = cols[check_b].scan_sum()
    assert got.count == n, "fused SUM count mismatch"
    assert (got.sum - int(dec.sum().item())) % (1 << 64) == 0, "fused SUM differs from the decoded vectors"
    from tests.cases import splitmix64
    head = splitmix64(np.arange(1 << 16, dtype=np.uint64) + np.uint64(SEED0 + check_b + rank * n))
    head &= np.uint64((1 << check_b) - 1) if check_b < 64 else np.uint64(2**64 - 1)
    assert np.array_equal(dec[: 1 << 16].cpu().numpy().view(np.uint64), head), "decoded values differ from the input"
    for b in widths:
        cols[b].uncompact()
    ctx.synchronize()

    logical_per_step = len(OPS) * raw_bytes * len(widths) * world
    value = logical_per_step / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: same five ops through the C ABI with host (pinned) buffers ----
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, ctx, cols, widths, counts, n, S, rank, world, dev)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        r = cpu_sweep(widths, args.cpu_log2_values, threads, steps=1, warmup=0)
        cpu = {"value": round(r["value"], 3), "unit": "GB/s", "cores": threads, "kind": r["kind"],
               "sample": r["sample"], "ops_GBs": r["ops_GBs"]}

    # ---- the other workload shapes of BASELINE.json (configs 1, 3, 4 on one GPU; 5 = sharded Q6 over NCCL) ----
    for c in cols.values():
        c.close()
    cols = {}
    configs = run_configs(args, ctx, rank, world, peak)

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": shared_config(args),
            "parallelism": f"row groups sharded, {world} rank(s), NCCL all-gather of 64 aggregates/step",
            "segments_per_column": int(len(counts)),
            "ops": op_summary, "configs": configs, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            # the drop-in path: SuccinctGpuFun's vtable slots driven like ColumnData::ScanVector / FetchRow (C++ host shim)
            "e2e_vtable": (configs or {}).pop("e2e_vtable", None),
            "gpu_launches": int(launches), "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
        if args.detail:
            with open(args.detail, "w") as f:
                json.dump(detail, f)
    for c in cols.values():
        c.close()
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_configs(args, ctx, rank, world, peak):
    wanted = [x for x in args.configs.split(",") if x]
    if not wanted:
        return None
    from tools import workloads
    threads = os.cpu_count() or 1
    out = {}

    def leg(key, fn, needs_single):
        if needs_single and world > 1:
            return  # one-GPU workloads: measured by the N = 1 run
        sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
        if rank == 0:
            sampler.start()
        t0 = time.time()
        try:
            res = fn()
        except Exception as ex:  # a failing leg must not take the headline line with it, but it must be visible
            res = {"error": f"{type(ex).__name__}: {ex}"}
        if rank == 0:
            res["clocks"] = sampler.stop()
            res["wall_s"] = round(time.time() - t0, 1)
            out[key] = res

    if "1" in wanted:
        leg("1_zipf_point_filters", lambda: workloads.config1_zipf(ctx, peak, threads), True)
    if "3" in wanted:
        leg("3_zipf_skew_sweep", lambda: workloads.config3_skews(ctx, peak, threads), True)
    if "4" in wanted:
        leg("4_tpch_sf10", lambda: workloads.config4_tpch(ctx, peak, threads), True)
    if "codecs" in wanted:
        leg("f4_neighbour_codecs", lambda: workloads.neighbour_codecs(ctx, peak, threads), True)
    if "vtable" in wanted:
        leg("e2e_vtable", lambda: workloads.vtable_dropin(threads), True)
    if "5" in wanted:
        leg("5_tpch_sf100_sharded", lambda: workloads.config5_tpch_sharded(ctx, peak, rank, world), False)
    return out if rank == 0 else None


def run_e2e(args, ctx, cols, widths, counts, n, S, rank, world, dev):
    import torch
    import torch.distributed as dist
    from duckdb_adaptive_compression_b200 import numa
    raw_bytes = n * S
    # threads + pinned staging of this rank onto the NUMA node of its GPU (restored at the end of the leg: the CPU
    # baseline uses all host cores)
    affinity_before = os.sched_getaffinity(0)
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    placement = numa.bind_to_gpu_node(local_rank, local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", str(world))))
    # pinned budget: raw per width + a packed and a flat buffer per host thread; keep half of the host memory free
    avail = host_mem_available_bytes()
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    budget = int(avail * 0.45 / max(local_world, 1))
    max_w = max(0, (budget - 8 * raw_bytes - (64 << 20)) // raw_bytes)
    if world > 1:
        t = torch.tensor([max_w], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        max_w = int(t.item())
    if max_w < 1:
        os.sched_setaffinity(0, affinity_before)
        return {"value": None, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": f"not enough host memory for pinned staging ({avail >> 30} GiB available)"}
    if max_w >= len(widths):
        e_widths = list(widths)
    else:  # evenly spaced subset of the sweep, always containing the extremes
        idx = sorted(set(int(round(i * (len(widths) - 1) / max(max_w - 1, 1))) for i in range(max_w)))
        e_widths = [widths[i] for i in idx]
    h_raw = {}
    for b in e_widths:
        arr = ctx.pinned_array(n, np.int64)
        # fill the host buffer with the same synthetic column (device generator -> D2H, outside the timing)
        cols[b].decode(out=arr)
        h_raw[b] = arr
    # A few host threads, each with its own context (= its own stream) and staging buffers, take alternate widths:
    # the H2D copy of one column overlaps the D2H copies of the other (PCIe is full duplex), exactly what a
    # multi-threaded DuckDB scan would do through the C ABI.  ctypes releases the GIL during the calls.
    import duckdb_adaptive_compression_b200 as adacom
    n_workers = 4
    workers = []
    for wi in range(n_workers):
        wctx = ctx if wi == 0 else adacom.Context(ctx.device)
        workers.append(dict(ctx=wctx, words=wctx.pinned_array(n + 4096 * 16, np.uint64),
                            flat=wctx.pinned_array(n, np.int64)))
    ctx.synchronize()

    def worker_run(wk, my_widths, out):
        wctx, h2d, d2h, agg = wk["ctx"], 0, 0, None
        for b in my_widths:
            col, metas, total = wctx.compress_host(h_raw[b], counts, words_out=wk["words"])  # H2D raw, D2H packed+meta
            col.decode(out=wk["flat"])                                                       # D2H decoded vectors
            agg = col.scan_sum()                                                             # D2H aggregate
            col.uncompact()
            col.close()
            h2d += raw_bytes
            d2h += total * 8 + len(metas) * 64 + raw_bytes + 32
        out.append((h2d, d2h, agg))

    def e2e_step():
        outs = [[] for _ in workers]
        threads = [threading.Thread(target=worker_run, args=(wk, e_widths[wi::n_workers], outs[wi]))
                   for wi, wk in enumerate(workers)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        res = [o[0] for o in outs if o]
        assert len(res) == len([1 for wi in range(n_workers) if e_widths[wi::n_workers]]), "an e2e worker failed"
        return sum(r[0] for r in res), sum(r[1] for r in res), res[0][2]

    def barrier():
        ctx.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    e2e_step()  # warm-up
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        h2d, d2h, agg = e2e_step()
    barrier()
    dt = (time.perf_counter() - t0) / args.e2e_steps
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    assert agg.count == n
    logical = len(OPS) * raw_bytes * len(e_widths) * world
    for arr in h_raw.values():
        ctx.free_pinned(arr)
    for wk in workers:
        wk["ctx"].free_pinned(wk["words"])
        wk["ctx"].free_pinned(wk["flat"])
        if wk["ctx"] is not ctx:
            wk["ctx"].close()
    try:
        os.sched_setaffinity(0, affinity_before)
    except OSError:
        pass
    return {"value": round(logical / dt / 1e9, 2), "unit": "GB/s", "h2d_bytes_per_step": int(h2d), "placement": placement,
            "d2h_bytes_per_step": int(d2h), "steps": args.e2e_steps, "warmup": 1, "ms_per_step": round(dt * 1e3, 2),
            "widths": e_widths if len(e_widths) < len(widths) else args.widths,
            "host_threads": n_workers,
            "note": "host (pinned) buffers through the C ABI: raw H2D, packed+meta D2H, decoded vectors D2H, "
                    "aggregate D2H; host threads take alternate widths so H2D and D2H overlap; PCIe-bound"}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
