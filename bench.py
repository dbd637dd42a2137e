#!/usr/bin/env python
"""Benchmark of the offer-evaluation hot path (BASELINE.json: offer x pod evaluations/sec, placement-cycle latency).

  python bench.py --gpus 1 --steps K --warmup W                 our arm (CUDA engine through the C ABI)
  python bench.py --impl reference --gpus N --steps K --warmup W  the reference path on the host cores (CPU oracle)

A "step" is one pass of the hot path over one batch: all offers x all requirements of the workload.
  value : N*M / t(K0 + K1 + K1x) with the tables resident in HBM (device time, CUDA events on the engine's stream)
  e2e   : N*M / t(one placement cycle through the C ABI from pinned HOST tables: H2D of every table, K0, K1, K1x, K2,
          D2H of the decisions)
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "offer x pod-instance evaluations/sec"
UNIT = "evaluations/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--workload", default="c5", choices=["c5", "c4", "c5-hostname", "c5-mount", "c5-redeploy", "c5-one-region"],
                    help="c5 = the headline configuration; c5-* = same size, shapes that leave its fast paths (DESIGN.md section 6)")
    ap.add_argument("--no-oracle-check", action="store_true", help="skip the decisions-vs-oracle check and the factored CPU leg (about 30 s)")
    ap.add_argument("--no-ncu", action="store_true", help="do not run the two-launch ncu pass that measures roofline.traffic and roofline.issue")
    ap.add_argument("--wire", default="packed", choices=["packed", "plain"],
                    help="offer table as oe_load_offers_packed takes it (narrow ids, fp32-exact scalars, `ports` as ids of distinct range lists) or full width")
    ap.add_argument("--probe", default=None, help=argparse.SUPPRESS)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debugging only; invalid as a bench number)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    return ap.parse_args()


def make_workload(args):
    from dcos_commons_b200 import synth
    if args.workload == "c5":
        return synth.config_c5(args.scale)
    if args.workload == "c4":
        return synth.config_c4(args.scale)
    return synth.config_c5_variant(args.workload[3:], args.scale)


def table_bytes(t, lo=0, hi=None):
    """Bytes of the encoded tables (what one cycle uploads)."""
    n = t.n_offers
    hi = n if hi is None else hi
    frac = (hi - lo) / max(n, 1)
    offer_cols = ("scalar", "scalar_present", "port_range", "mount_size", "mount_profile", "alloc_role", "agent_id", "hostname_id",
                  "zone_id", "region_id", "attr_id", "rr_key", "exec_id")
    other = ("op_begin", "ops", "rule_nodes", "pod_role", "port_excl_begin", "port_excl", "range_set_begin", "range_set", "counter_sets",
             "type_set_pod_type", "matcher_bits", "t_name_id", "t_pod_type_id", "t_pod_index", "t_agent_id", "t_hostname_id", "t_zone_id",
             "t_region_id", "t_attr_id", "t_rr_key", "t_filter_bits", "class_id", "r_pod_type_id", "r_pod_index", "launch_begin",
             "launch_name_id", "launch_filter_bits")
    b_off = sum(getattr(t, c).nbytes for c in offer_cols)
    b_other = sum(getattr(t, c).nbytes for c in other if getattr(t, c) is not None)
    return int(b_off * frac) + b_other, b_off // max(n, 1)


def algorithmic_bytes(t, n_local):
    """every offer row read once, every requirement once, one result bit per pair written once."""
    _, b_offer = table_bytes(t)
    b_class = (t.ops.nbytes + t.rule_nodes.nbytes + t.matcher_bits.nbytes + t.port_excl.nbytes) // max(t.n_classes, 1)
    d = t.dict_sizes
    ks = {0: d["n_hostnames"], 1: d["n_zones"], 2: d["n_regions"], 3: d["n_attrs"], 4: d["n_rr_values"]}
    counters = sum(ks[int(c["keyspace"])] * 4 for c in t.counter_sets)
    M, C = t.n_reqs, t.n_classes
    words = (n_local + 31) // 32
    k1 = n_local * b_offer + C * b_class + counters + C * words * 4      # class-factored feasibility
    k1x = C * words * 4 + M * 4 + M * words * 4                            # expansion: read class rows + req->row, write matrix
    return dict(k1=k1, k1x=k1x, total=n_local * b_offer + M * 4 + C * b_class + counters + M * words * 4, b_offer=b_offer, b_class=b_class)


class ClockSampler:
    """Samples SM clock and throttle reasons DURING the timed region (NVML at ~1 kHz; nvidia-smi when NVML is missing)."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self.thread = None

    def _nvml_init(self):
        import pynvml as nv
        nv.nvmlInit()
        self._nv = nv
        self._h = nv.nvmlDeviceGetHandleByIndex(self.index)
        self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM))
        self._get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons

    def _nvml_sample(self):
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        self.sm.append(float(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM)))
        r = int(self._get_reasons(self._h))
        for name, bit in bits.items():
            if r & bit:
                self.reasons.add(name)

    def _nvml_loop(self):
        while not self._stop.is_set():
            self._nvml_sample()
            time.sleep(0.0002)

    def _smi_loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout
            except Exception:
                return
            r = [x.strip() for x in out.strip().split(",")]
            if len(r) >= 9:
                try:
                    self.sm.append(float(r[1]))
                    self.max_mhz = float(r[2])
                except ValueError:
                    pass
                for k, name in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
                    if r[5 + k].lower() == "active":
                        self.reasons.add(name)

    def start(self):
        try:
            self._nvml_init()   # before the timed region starts: a short region (N = 8: a few ms) must still be sampled
            target = self._nvml_loop
        except Exception:
            target = self._smi_loop
        self.thread = threading.Thread(target=target, daemon=True)
        self.thread.start()

    def stop(self):
        self._stop.set()
        if self.thread:
            self.thread.join(timeout=6)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.sm)}


def pin_tables(t):
    """Copies every column into pinned host memory (torch is only the allocator here)."""
    import copy
    import torch
    p = copy.copy(t)
    p._keep = []
    for name, val in vars(t).items():
        if isinstance(val, np.ndarray) and val.size:
            if val.dtype.fields is not None:  # structured dtypes: pin as raw bytes
                raw = torch.from_numpy(val.view(np.uint8).copy()).pin_memory()
                setattr(p, name, raw.numpy().view(val.dtype).reshape(val.shape))
            else:
                setattr(p, name, torch.from_numpy(val.copy()).pin_memory().numpy())
    return p


def cpu_baseline_sample(t, seconds, n_sample_offers=1000):
    """Reference path (CPU oracle, 1 thread) on a bounded sample of the same workload: whole requirement rows over the
    first `n_sample_offers` offers, until the time budget is used."""
    from oracle_binding import Oracle
    import copy
    n = min(n_sample_offers, t.n_offers)
    orc = Oracle()
    orc.tables = t
    orc.load_programs(t)
    orc.load_tasks(t)
    orc.load_offers(t, 0, n)
    orc.load_requirements(t)
    t0 = time.perf_counter()
    rows = 0
    while rows < t.n_reqs and (time.perf_counter() - t0 < seconds or rows == 0):
        orc.feasibility_rows(rows, rows + 1)
        rows += 1
    dt = time.perf_counter() - t0
    orc.close()
    return {"value": rows * n / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{rows} requirement rows x first {n} offers of the workload ({t.n_tasks} deployed tasks walked per rule, as the reference does), "
                      f"{dt:.1f} s, CPU oracle oracle/oe_oracle.cpp -O2"}



def oracle_check_and_factored_cpu(t, offer_rows, full_decisions=None):
    """The decisions of the last timed cycle against the CPU oracle's own cycle on the same tables (cached mode: tallies built
    once per cycle, pinned to the literal mode by tests/test_oracle_modes.py), and — from the same run — what the host cores do
    when THEY use the engine's factoring (one bit row per class instead of per requirement): NOT reference behaviour, reported so
    that the share of the speed-up that is algorithm and the share that is GPU can be told apart."""
    from oracle_binding import Oracle
    cores = os.cpu_count() or 1
    orc = Oracle(cached=True, threads=cores)
    orc.load(t)
    sample_classes = list(range(0, t.n_classes, max(1, t.n_classes // 4)))[:4]
    t0 = time.perf_counter()
    for c in sample_classes:
        j = int(np.flatnonzero(t.class_id == c)[0])
        orc.feasibility_row(j)
    t_rows = (time.perf_counter() - t0) * t.n_classes / len(sample_classes)
    t0 = time.perf_counter()
    d_cpu, st = orc.evaluate_cycle()
    t_cycle = time.perf_counter() - t0
    orc.close()
    same_offers = bool(np.array_equal(d_cpu["offer_row"], offer_rows))
    same_all = None if full_decisions is None else bool(d_cpu.tobytes() == full_decisions.tobytes())
    factored = {"value": t.n_offers * t.n_reqs / (t_rows + t_cycle), "unit": UNIT, "cores": cores, "kind": "port, factored (not reference behaviour)",
                "sample": f"{len(sample_classes)} of {t.n_classes} class rows x all {t.n_offers} offers on {cores} threads ({t_rows:.1f} s for all rows, extrapolated) "
                          f"+ the whole ordered cycle on 1 thread ({t_cycle:.1f} s, {st['n_pairs']} pairs): tallies built once, one bit row per class"}
    return same_offers, same_all, factored


def ncu_probe(args, shard):
    """DRAM bytes and warp instructions per launch of K1 and K1x, measured now: a second process runs two passes of the same
    workload under `ncu` (nothing timed there).  None when ncu is not usable on this box."""
    import csv
    import shutil
    import tempfile
    ncu = shutil.which("ncu") or "/usr/local/cuda/bin/ncu"
    if not os.path.exists(ncu):
        return None, "ncu not found"
    log = os.path.join(tempfile.mkdtemp(prefix="oe_ncu_"), "probe.csv")
    cmd = [ncu, "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum", "--clock-control", "none",
           "-k", "regex:k_expand_matrix|k_class_feasibility", "-c", "4", "--csv", "--log-file", log,
           sys.executable, os.path.abspath(__file__), "--probe", shard, "--workload", args.workload, "--scale", str(args.scale), "--wire", args.wire]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")})
    except Exception as ex:  # noqa: BLE001
        return None, f"ncu pass failed: {ex}"
    if r.returncode != 0 or not os.path.exists(log):
        return None, f"ncu pass exited {r.returncode}: {(r.stderr or r.stdout)[-200:]}"
    rows = [ln for ln in open(log) if ln.startswith('"')]
    acc = {}
    for rec in csv.DictReader(rows):
        name = "k_expand_matrix" if "k_expand_matrix" in rec["Kernel Name"] else "k_class_feasibility"
        val = float(rec["Metric Value"].replace(",", ""))
        unit = rec.get("Metric Unit", "")
        if rec["Metric Name"].startswith("dram__bytes"):
            val *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        acc.setdefault(name, {}).setdefault(rec["Metric Name"], []).append(val)
    out = {k: {m: sum(v) / len(v) for m, v in d.items()} for k, d in acc.items()}
    return (out or None), ("ncu pass of this run" if out else "ncu pass produced no rows")


def run_probe(args):
    """Child of ncu_probe: two passes of K0 + K1 + K1x over the shard `lo,hi,shard_rows,rank,world` (no timing, no output)."""
    from dcos_commons_b200 import Engine
    lo, hi, shard_rows, rank, world = (int(x) for x in args.probe.split(","))
    t = make_workload(args).tables
    from dcos_commons_b200 import abi
    eng = Engine(device=0, rank=rank, world=world)
    eng.tables = t
    eng.load_programs(t)
    eng.load_tasks(t)
    if args.wire == "packed":
        eng.load_offers_packed(abi.PackedOffers(t, lo, hi), shard_rows)
    else:
        eng.load_offers(t, lo, hi, shard_rows)
    eng.load_requirements(t)
    for _ in range(2):
        eng.feasibility(fetch=False)
    eng.close()

# ----------------------------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own CPU path for the same metric: the CPU oracle (the JVM reference cannot run on this box), one
    process per host core, each evaluating whole requirement rows over a sample of the workload's offers."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cl = make_workload(args)
    t = cl.tables
    cores = os.cpu_count() or 1
    global _REF_T
    _REF_T = t
    # calibrate the sample so that one step takes a few seconds
    probe = cpu_baseline_sample(t, 1.0, n_sample_offers=200)
    per_core = probe["value"]
    n = int(min(t.n_offers, max(200, per_core * 3.0)))  # ~3 s per row per core
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_ref_init, initargs=(n,)) as pool:
        times = []
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pool.map(_ref_row, [(s * cores + k) % t.n_reqs for k in range(cores)])
            if s >= args.warmup:
                times.append(time.perf_counter() - t0)
    total = sum(times)
    value = args.steps * cores * n / total
    sample = f"each step: {cores} requirement rows (one per process) x first {n} offers; {t.n_tasks} deployed tasks"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64+int32",
            "data": "synthetic", "config": {"workload": cl.name, "offers": t.n_offers, "requirements": t.n_reqs, "classes": t.n_classes,
                                            "deployed_tasks": t.n_tasks},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


_REF_T = None
_REF_ORC = None


def _ref_init(n):
    global _REF_ORC
    from oracle_binding import Oracle
    t = _REF_T
    o = Oracle()
    o.tables = t
    o.load_programs(t)
    o.load_tasks(t)
    o.load_offers(t, 0, n)
    o.load_requirements(t)
    _REF_ORC = o


def _ref_row(j):
    _REF_ORC.feasibility_rows(j, j + 1)
    return 0


# ----------------------------------------------------------------------------------------------------------------------
def run_engine(args):
    import torch
    from dcos_commons_b200 import Engine, abi, sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)

    cl = make_workload(args)
    t = cl.tables
    N, M = t.n_offers, t.n_reqs
    shard_rows, bounds = sharded.shard_plan(N, world)
    lo, hi = bounds[rank]
    n_local = hi - lo

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- resident-in-HBM arm -------------------------------------------------------------------------------------------
    packed = abi.PackedOffers(t, lo, hi) if args.wire == "packed" else None

    def load_all(e, tabs, pk):
        e.tables = tabs
        e.load_programs(tabs)
        e.load_tasks(tabs)
        if pk is not None:
            e.load_offers_packed(pk, shard_rows)
        else:
            e.load_offers(tabs, lo, hi, shard_rows)
        e.load_requirements(tabs)

    eng = Engine(device=local_rank, rank=rank, world=world)
    load_all(eng, t, packed)
    value_exchange = None
    if world > 1:   # the pass that is timed publishes its class rows to every shard, as the sharded cycle's pass does
        ok = os.environ.get("OE_BENCH_EXCHANGE", "peer") == "peer" and sharded.attach_peers(eng, dist, max_rows=max(64, 2 * eng.factored_rows()))
        value_exchange = "peer stores inside the feasibility kernel" if ok else "none (no peer access): rows stay local"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    ext = torch.cuda.ExternalStream(eng.cuda_stream())

    def flush_l2():
        with torch.cuda.stream(ext):
            flush.zero_()

    for _ in range(args.warmup):
        flush_l2()
        eng.feasibility(fetch=False)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_steps, k0, k1, k1x, launches = [], 0.0, 0.0, 0.0, 0
    for _ in range(args.steps):
        flush_l2()
        _, st = eng.feasibility(fetch=False)   # synchronises the engine stream; st holds CUDA-event times of this step
        ms_steps.append(st["ms_counters"] + st["ms_feasibility"] + st["ms_expand"])
        k0 += st["ms_counters"]; k1 += st["ms_feasibility"]; k1x += st["ms_expand"]
        launches += st["kernel_launches"]
    barrier()
    clocks = sampler.stop()
    dev_ms = sum(ms_steps)
    if dist is not None:
        tmax = torch.tensor([dev_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dev_ms = float(tmax.item())
    value = args.steps * N * M / (dev_ms * 1e-3)

    # ---- end-to-end arm: pinned host tables -> C ABI -> decisions in pinned host memory ------------------------------------
    pt = pin_tables(t.cut_offers(lo, hi) if world > 1 else t)   # each rank keeps only its shard of the offers, pinned
    ppk = abi.PackedOffers(t, lo, hi).pin() if args.wire == "packed" else None   # ... in the wire form the encoder emits
    out = torch.empty(M * abi.DECISION_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    # the tables stay alive in pinned memory for the whole cycle, so the loads may return before their copies have landed
    e2e_matrix = os.environ.get("OE_BENCH_E2E_MATRIX", "1") != "0"     # 0: the timed cycle asks for the decisions only (diagnostics)
    e2e_flags = abi.OE_CFG_ASYNC_LOADS | (abi.OE_CFG_MATERIALIZE_MATRIX if e2e_matrix else 0)
    eng2 = Engine(device=local_rank, rank=rank, world=world, flags=e2e_flags)

    host_ms = {}

    def timed(name, fn):
        t0 = time.perf_counter()
        r = fn()
        host_ms[name] = host_ms.get(name, 0.0) + 1e3 * (time.perf_counter() - t0)
        return r

    exchange = None
    merged_on_rank0 = False

    def cycle():
        eng2.tables = pt
        timed("load_programs", lambda: eng2.load_programs(pt))
        timed("load_tasks", lambda: eng2.load_tasks(pt))
        timed("load_offers", (lambda: eng2.load_offers_packed(ppk, shard_rows)) if ppk is not None else (lambda: eng2.load_offers(pt, lo, hi, shard_rows)))
        timed("load_requirements", lambda: eng2.load_requirements(pt))
        if world > 1 and exchange == "nccl all-gather":
            eng2.feasibility(fetch=False)                 # K0 + K1 + K1x on this rank's shard
            sharded.gather_class_bits(eng2, dist)         # the one exchange step: class rows of every shard
        # peers attached: the pass publishes its rows into every rank's exchange region itself, this call is the whole cycle;
        # every rank pushes its winners' ports / volumes into every region, and only rank 0 fetches the merged table
        if world > 1 and merged_on_rank0:
            st = timed("evaluate_cycle", lambda: eng2.evaluate_cycle_into(None))
            if rank == 0:
                timed("gather_decisions", lambda: eng2.gather_decisions_into(out.data_ptr()))
            return st
        return timed("evaluate_cycle", lambda: eng2.evaluate_cycle_into(out.data_ptr()))   # [K0 + K1 (+ K1x beside)] K3 + K2 (replayed identically on every rank) + D2H

    if world > 1:
        # the exchange regions are wired up once, outside the cycles (a scheduler does it at start-up): needs the offers' shape
        load_all(eng2, pt, ppk)
        attached = os.environ.get("OE_BENCH_EXCHANGE", "peer") == "peer" and sharded.attach_peers(eng2, dist, max_rows=max(64, 2 * eng2.factored_rows()))
        exchange = "peer stores from the feasibility kernel (NVLink), flags, device-side wait" if attached else "nccl all-gather"
        merged_on_rank0 = bool(attached)
        if exchange == "nccl all-gather":
            eng2.close()
            eng2 = Engine(device=local_rank, rank=rank, world=world, flags=e2e_flags)

    cyc_stats = None
    for _ in range(max(1, min(args.warmup, 3))):
        cyc_stats = cycle()
    barrier()
    host_ms.clear()
    e2e_steps = max(1, args.steps)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        cyc_stats = cycle()
    barrier()
    e2e_s = time.perf_counter() - t0
    # what the last timed cycle decided: every rank's digest must be the same one (the reference has ONE list of offers,
    # S/scheduler/plan/PlanScheduler.java:56-67), and the same at every N
    import hashlib
    if world > 1 and merged_on_rank0 and rank != 0:
        eng2.gather_decisions_into(out.data_ptr())     # outside the timing: every rank's own table, for the agreement check below
    decisions = np.frombuffer(out.numpy().tobytes(), dtype=abi.DECISION_DTYPE)
    offer_rows = np.ascontiguousarray(decisions["offer_row"])
    digest = hashlib.sha256(offer_rows.tobytes()).hexdigest()
    digest_all_fields = hashlib.sha256(decisions.tobytes()).hexdigest()
    digests = [digest + digest_all_fields if merged_on_rank0 else digest]
    if dist is not None:
        digests = [None] * world
        dist.all_gather_object(digests, digests_mine := (digest + digest_all_fields if merged_on_rank0 else digest))
    if dist is not None:
        tmax = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        e2e_s = float(tmax.item())
    h2d, _ = table_bytes(t, lo, hi)
    if ppk is not None:
        h2d = table_bytes(t, 0, 0)[0] + ppk.nbytes     # programs, tasks, requirements at full width + the packed offer columns
    d2h = M * abi.DECISION_DTYPE.itemsize + 4
    e2e = {"value": e2e_steps * N * M / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "cycle_latency_ms": 1e3 * e2e_s / e2e_steps, "steps": e2e_steps,
           "device_ms": {k: cyc_stats[k] for k in ("ms_counters", "ms_feasibility", "ms_expand", "ms_commit", "ms_d2h")},
           "host_ms_per_call": {k: v / e2e_steps for k, v in host_ms.items()}, "placed": cyc_stats["n_placed"],
           "decisions_sha256": digest, "ranks_agree": len(set(digests)) == 1, "matrix_written": e2e_matrix}
    if len(set(digests)) != 1:
        raise SystemExit(f"bench: the ranks disagree on the placement decisions: {digests}")
    if exchange:
        e2e["exchange"] = exchange

    # the same cycle when the caller only wants the decisions (what the scheduler adapter asks for): no matrix is written
    if world == 1:
        eng3 = Engine(device=local_rank, flags=abi.OE_CFG_ASYNC_LOADS)

        def cycle_decisions_only():
            load_all(eng3, pt, ppk)
            return eng3.evaluate_cycle_into(out.data_ptr())
        for _ in range(2):
            cycle_decisions_only()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            st3 = cycle_decisions_only()
        e2e["decisions_only"] = {"cycle_latency_ms": 1e3 * (time.perf_counter() - t0) / e2e_steps,
                                 "device_ms": {k: st3[k] for k in ("ms_counters", "ms_feasibility", "ms_commit", "ms_d2h")},
                                 "note": "same tables and copies, OE_CFG_MATERIALIZE_MATRIX off: K0 + K1 + commit, the N x M matrix stays factored"}
        # steady state (INTEGRATION.md "Offers that change between cycles"): the table stays resident, 1 % of the offers changed
        # since the last cycle and arrive through oe_update_offers; tasks unchanged; new requirements; decisions only
        import ctypes as C
        rng = np.random.Generator(np.random.PCG64(7))
        changed = np.sort(rng.choice(N, max(1, N // 100), replace=False)).astype(np.int64)
        small = pt.c_offer_rows(changed)               # the adapter's encoding of the changed offers, prepared outside the timing
        upd = eng3.b.fn("update_offers")
        upd.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]

        def cycle_steady():
            eng3._check(upd(eng3.h, changed.ctypes.data, len(changed), C.byref(small)))
            eng3.load_requirements(pt)
            return eng3.evaluate_cycle_into(out.data_ptr())
        for _ in range(2):
            cycle_steady()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            st4 = cycle_steady()
        e2e["steady_state"] = {"cycle_latency_ms": 1e3 * (time.perf_counter() - t0) / e2e_steps, "changed_offers": int(len(changed)),
                               "device_ms": {k: st4[k] for k in ("ms_counters", "ms_feasibility", "ms_commit", "ms_d2h")},
                               "note": "resident offer table, 1 % of the rows replaced through oe_update_offers, requirements reloaded, decisions only"}
        eng3.close()

    if dist is not None:   # nothing below is collective: rank 0 alone measures the CPU legs and the ncu pass
        barrier()
        dist.destroy_process_group()
        dist = None
    if rank != 0:
        return

    # ---- roofline of the dominant kernel -------------------------------------------------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    ab = algorithmic_bytes(t, n_local)
    k1_ms, k1x_ms = k1 / args.steps, k1x / args.steps
    kernels = [
        {"kernel": "k_expand_matrix", "ms": k1x_ms, "algorithmic_bytes": ab["k1x"], "achieved_gbs": ab["k1x"] / (k1x_ms * 1e-3) / 1e9 if k1x_ms else None},
        {"kernel": "k_class_feasibility", "ms": k1_ms, "algorithmic_bytes": ab["k1"], "achieved_gbs": ab["k1"] / (k1_ms * 1e-3) / 1e9 if k1_ms else None},
        {"kernel": "k_count_tasks+k_rr_summary", "ms": k0 / args.steps},
    ]
    dom = max(kernels[:2], key=lambda k: k["ms"])
    # DRAM bytes and warp instructions per launch: measured by a two-pass ncu run of this same workload and shard right now;
    # when ncu cannot run here, the committed capture of the headline configuration (one GPU) or null, and the line says which
    probe, probe_src = (None, "--no-ncu") if args.no_ncu else ncu_probe(args, f"{lo},{hi},{shard_rows},{rank},{world}")
    traffic = insts = k1_ncu_ns = None
    if probe and dom["kernel"] in probe:
        traffic = int(probe[dom["kernel"]].get("dram__bytes_read.sum", 0) + probe[dom["kernel"]].get("dram__bytes_write.sum", 0))
        traffic_src = probe_src
    else:
        traffic_path = os.path.join(ROOT, "profiles", "kernel_traffic.json")
        traffic_src = f"null: {probe_src}"
        if world == 1 and args.workload == "c5" and args.scale == 1.0 and os.path.exists(traffic_path):
            kt = json.load(open(traffic_path)).get(dom["kernel"])
            if kt:
                traffic = kt["dram_bytes_read"] + kt["dram_bytes_write"]
                traffic_src = f"committed capture profiles/kernel_traffic.json ({probe_src})"
    if probe and "k_class_feasibility" in probe:
        insts = probe["k_class_feasibility"].get("smsp__inst_executed.sum")
        k1_ncu_ns = probe["k_class_feasibility"].get("gpu__time_duration.sum")
    # the feasibility kernel is bound by issue slots, not by HBM — 148 SMs x 4 schedulers x SM clock
    sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
    rows_evaluated = eng.factored_rows()
    issue = {"kernel": "k_class_feasibility", "warp_inst": insts, "inst_per_pair": insts * 32 / max(n_local * rows_evaluated, 1) if insts else None,
             "issue_slots_per_s": 148 * 4 * sm_hz, "issue_frac": insts / (148 * 4 * sm_hz * k1_ms * 1e-3) if insts and k1_ms else None,
             "pairs": n_local * rows_evaluated, "ms": k1_ms, "hbm_frac": (ab["k1"] / (k1_ms * 1e-3) / 1e9 / peak) if k1_ms else None,
             "kernel_ms_under_ncu": k1_ncu_ns * 1e-6 if k1_ncu_ns else None,
             "issue_frac_kernel_alone": insts / (148 * 4 * sm_hz * k1_ncu_ns * 1e-9) if insts and k1_ncu_ns else None,
             "source": probe_src if insts else f"null: {probe_src}",
             "note": "pairs = offers x factored bit rows (classes + re-deployed pods); inst_per_pair = warp instructions x 32 lanes / pairs; "
                     "ms = live CUDA-event time of the feasibility phase (the rule / port / transpose table kernels and their launch gaps included), "
                     "kernel_ms_under_ncu = the kernel's own duration in the ncu pass (serialised, cold cache)"}
    roofline = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved_gbs"], "peak": peak, "unit": "GB/s",
                "frac": dom["achieved_gbs"] / peak if dom["achieved_gbs"] else None, "traffic": traffic,
                "traffic_source": traffic_src,
                "algorithmic_bytes": dom["algorithmic_bytes"], "peak_source": peak_src,
                "pass_achieved": ab["total"] / (dev_ms / args.steps * 1e-3) / 1e9, "pass_frac": ab["total"] / (dev_ms / args.steps * 1e-3) / 1e9 / peak,
                "pass_note": "pass_* = SURVEY.md §8(d) bytes_K1 (un-factored: one bit per pair written) over K0+K1+K1x device time",
                "bytes_per_eval": ab["total"] / max(n_local * M, 1), "kernels": kernels, "issue": issue}
    if value_exchange:
        roofline["pass_exchange"] = value_exchange

    cpu = cpu_baseline_sample(t, args.cpu_seconds) if world == 1 else None
    factored = None
    if not args.no_oracle_check:
        same_offers, same_all, factored = oracle_check_and_factored_cpu(t, offer_rows, decisions if (world == 1 or merged_on_rank0) else None)
        e2e["decisions_match_oracle"] = same_offers
        if same_all is not None:
            e2e["decisions_match_oracle_all_fields"] = same_all     # dynamic ports, MOUNT slots, fresh reservation ids too
        if not same_offers or same_all is False:
            raise SystemExit("bench: the engine's placement decisions differ from the CPU oracle's")
    digests_path = os.path.join(ROOT, "profiles", "decisions_digests.json")
    if os.path.exists(digests_path) and args.scale == 1.0:
        committed = json.load(open(digests_path)).get(cl.name)
        e2e["decisions_sha256_committed"] = committed
        if committed and committed != digest:
            raise SystemExit(f"bench: decisions digest {digest} differs from the committed one {committed} for {cl.name}")

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64+int32", "data": "synthetic",
            "config": {"workload": cl.name, "offers": N, "requirements": M, "classes": t.n_classes, "deployed_tasks": t.n_tasks,
                       "offers_per_gpu": n_local, "wire": args.wire, "parallelism": f"offers row-sharded x{world}" if world > 1 else "single GPU",
                       "l2": "256 MiB memset between timed steps flushes L2; the written matrix (N*M/8 bytes per step) is far larger than L2"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if factored is not None:
        line["cpu_baseline_factored"] = factored
    print(json.dumps(line))


def main():
    args = parse_args()
    # exactly ONE line may reach stdout: libraries that print there (NCCL's version banner) are sent to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.probe:
        run_probe(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_engine(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
