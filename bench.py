"""bench.py — QMMX Monte Carlo hot path on B200: simulated path-steps/s (390-step paths).

    python bench.py --gpus 1 --steps 20 --warmup 3            # our arm (libqmmx_mc.so, sm_100a)
    python bench.py --impl reference --gpus 1 --steps 3 ...   # the reference's CPU path (oracle port), host cores
    torchrun --nproc-per-node N ... bench.py --gpus N ...     # one rank per GPU, paths sharded, one NCCL all-gather + merge kernel

A "step" = one pass of the engine scan over one batch of resident synthetic bars (BASELINE.json configs[1]:
1M GBM paths x 390 one-minute closes per GPU, the 30-level Blue/Orange/Black/Teal x Solid/Dashed fixture, sklearn-
form LR entry filter enabled) + the reduction of PnL moments and reason-code histograms (+ the merge over ranks, N>1).
Prints ONE JSON line on rank 0.  The same line carries, outside the timed region of the headline: `c3` (BASELINE
configs[2], the target config: 100M Philox paths x 390 STRONG-scaled over the N GPUs, fused generate+scan, one merge at
the end, with an aggregate digest that must not depend on N), `c4`, `c5f`, `mc` (configs[3], configs[4], the
reference's own jitter Monte Carlo), in-run equality checks against the C oracle, the Python/NumPy CPU baselines and a
pinned host-to-device copy probe.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from qmmx_monolithic_monte_carlo_b200 import abi, fixtures as F  # noqa: E402

METRIC = "simulated path-steps/sec (390-step paths)"
UNIT = "path-steps/s"
STEPS_PER_PATH = 390


def workload(name: str):
    if name == "c2":
        return dict(workload="C2: 1M host GBM paths x 390 closes per GPU, 30 levels (4 colours x solid/dashed), "
                             "LR entry filter on, close-only exits",
                    paths=1_000_000, levels=F.levels_30(), lr=True)
    if name == "c1":
        return dict(workload="C1: 10k host GBM paths x 390 closes, one Blue/Solid level set", paths=10_000,
                    levels=F.levels_blue_solid_4(), lr=False)
    raise SystemExit(f"unknown workload {name}")


def lr_weights():
    """Synthetic sklearn-form LR (no trained model ships with the reference):
    x = [lvl_type(=1), |level - stop|, touch_count, dir_up] (M:1457-1461).  Chosen so that the gate is active:
    first touches pass both ways, second touches pass only with the trend up, third touches fail."""
    return np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])     # coef[4] + intercept


def engine_params():
    return abi.make_params(contact_prox=0.10, stop_padding=0.35, tp_padding=0.25, q_min_prob=0.60, cooldown_s=120)


class ClockSampler(threading.Thread):
    """nvidia-smi style clock / throttle-reason sampling (NVML) during the timed region."""

    def __init__(self, index: int, period_s: float = 0.004):
        super().__init__(daemon=True)
        self.index, self.period, self.samples, self.reasons, self.max_mhz = index, period_s, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def agg_digest(results) -> str:
    """sha256 over every INTEGER aggregate of every combo (counters, both reason histograms, the 256-bin PnL histogram):
    identical for any number of ranks when the paths are the same (float64 sums depend on the summation order and stay
    out)."""
    h = hashlib.sha256()
    for r in results:
        a = r.agg
        ints = [a.n_paths, a.n_steps, a.n_trades, a.n_wins, a.n_open] + list(a.entry_hist) + list(a.exit_hist) + list(a.pnl_hist)
        h.update(np.array(ints, dtype=np.uint64).tobytes())
    return h.hexdigest()[:24]


def int_aggs_equal(a, b) -> bool:
    return (all(getattr(a, f) == getattr(b, f) for f in ("n_paths", "n_steps", "n_trades", "n_wins", "n_open")) and
            list(a.entry_hist) == list(b.entry_hist) and list(a.exit_hist) == list(b.exit_hist) and list(a.pnl_hist) == list(b.pnl_hist) and
            abs(a.sum_pnl - b.sum_pnl) <= 1e-9 * max(1.0, abs(b.sum_pnl)))


def shipped_capture():
    """binding_limit numbers come from the committed ncu capture of the shipped build, not from literals"""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_shipped_capture.json")) as f:
            return json.load(f)
    except Exception:
        return None


def bind_numa(local_rank: int, world: int):
    """One rank per GPU: pin the process (and so the first-touch placement of its host bars) to the NUMA node of its GPU;
    when the platform reports no node per GPU, spread the ranks over the nodes round-robin.  QMMX_NUMA=off disables it."""
    info = {"policy": os.environ.get("QMMX_NUMA", "auto")}
    if info["policy"] == "off" or not hasattr(os, "sched_setaffinity"):
        return info
    try:
        nodes = sorted(int(d[4:]) for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit())
        node = None
        try:
            import pynvml
            pynvml.nvmlInit()
            bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
            bus = bus.decode() if isinstance(bus, bytes) else bus
            with open("/sys/bus/pci/devices/%s/numa_node" % bus[-12:].lower()) as f:
                node = int(f.read().strip())
        except Exception:
            node = None
        info["gpu_node"] = node
        if info["policy"] == "spread" or node is None or node < 0 or node not in nodes:
            node = nodes[local_rank % len(nodes)] if (world > 1 and len(nodes) > 1) else None
        if node is None:
            return info
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            info.update(node=node, cpus=len(allowed))
    except Exception as e:           # never fatal: the binding is an optimisation of the host-side copy
        info["error"] = str(e)[:80]
    return info


def headline_config(wl, world, n_levels):
    """`config` of the headline line: the same dict in both arms (ours and --impl reference), so the driver sees one config"""
    return {"workload": wl["workload"], "paths_per_gpu": wl["paths"], "steps_per_path": STEPS_PER_PATH,
            "levels": n_levels, "lr_filter": bool(wl["lr"]), "bars": "fp32 close [step][path], resident in HBM",
            "l2": "inputs (1.56 GB per GPU) larger than the 126 MB L2; no flush needed",
            "policy_gate": "off (DISABLE_POLICY_GATE=1, as in the reference's recorded run, SURVEY B.2)",
            "parallelism": f"paths sharded over {world} GPU(s); per step one NCCL all-gather of the 2.3 KB aggregate "
                           "vector + the library's merge kernel"}


def cpu_port_run(close, price, solid, params, lr, threads):
    from oracle import c_oracle as CO
    t0 = time.perf_counter()
    o = CO.run_paths(close, price, solid, params, lr=lr, n_threads=threads)
    dt = time.perf_counter() - t0
    return o["agg"], dt


def run_reference(args, wl, rank):
    """Reference arm: the reference's CPU implementation of the path = the oracle's C port (the reference itself
    is pure Python with SQLite in the loop), all host threads, a bounded sample of the same workload per step."""
    if rank != 0:
        return
    from qmmx_monolithic_monte_carlo_b200 import engine as E
    from oracle import c_oracle as CO
    CO.build()
    rows, price, solid = E.sort_levels(wl["levels"])
    params = engine_params()
    lr = lr_weights() if wl["lr"] else None
    threads = os.cpu_count() or 1
    sample = min(wl["paths"], 262_144)
    close = F.host_gbm_closes(sample, STEPS_PER_PATH)
    for _ in range(max(1, args.warmup)):
        cpu_port_run(close[:, : sample // 8], price, solid, params, lr, threads)
    t = 0.0
    for _ in range(args.steps):
        _, dt = cpu_port_run(close, price, solid, params, lr, threads)
        t += dt
    ms = 1e3 * t / args.steps
    value = sample * STEPS_PER_PATH / (ms * 1e-3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": headline_config(wl, args.gpus, len(rows)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "host_cpu": cpu_model(),
                         "sample": f"{sample} of {wl['paths']} paths x {STEPS_PER_PATH} steps per step, oracle/qmmx_oracle.c, pthreads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def fp32_roofline(eng, units_per_s, flop_per_unit):
    """Second roofline of SURVEY 8(d): contract flop per path-step (12 scan-only, 24 fused GBM) x measured rate,
    against the FFMA-only rate measured on this device in this run (outside the timed region)."""
    peak = eng.measure_fp32_peak()
    ach = flop_per_unit * units_per_s / 1e12
    return {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
            "flop_per_path_step": flop_per_unit, "peak_source": "qmmx_measure_fp32_peak: FFMA-only kernel, this run",
            "note": "essential flop only; the engine is compare/select/table work, so the issue-slot use "
                    "(smsp__issue_active in profiles/) is the binding figure, not this fraction"}


def run_extra(args, name, rank, local_rank, world):
    """Secondary configs of BASELINE.json (not the driver's headline line): c2hl (C2 with OHLC bars and exits on
    high/low), c3 (fused Philox GBM + scan, 12.5M paths per GPU), c4 (256-combo stop/target x cooldown sweep over
    resident bars), c5 (block bootstrap of a 40-bar window into 5-day paths, per-day LR weights)."""
    import torch
    import torch.distributed as dist
    from qmmx_monolithic_monte_carlo_b200 import distributed as D, engine as E
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    eng = E.MCEngine(local_rank, args.lib)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); eng.set_stream(stream.cuda_stream)
    eng.set_levels(F.levels_30())
    g = abi.make_gbm()
    cfg = {}
    if name == "c3":
        P = args.paths or 12_500_000
        eng.set_params(engine_params()); eng.set_lr(lr_weights())
        units = P * STEPS_PER_PATH
        run = lambda: eng.run_fused_gbm_async(F.HOST_SEED, rank * P, P, STEPS_PER_PATH, g)
        cfg = dict(workload=f"C3: {P} Philox GBM paths x 390 per GPU, generated and scanned in registers (no bar traffic)",
                   bytes_per_step=0)
    elif name == "c2hl":
        P = args.paths or 1_000_000
        eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)); eng.set_lr(lr_weights())
        eng.gen_gbm(F.HOST_SEED, rank * P, P, STEPS_PER_PATH, abi.make_gbm(sub_steps=4))
        units = P * STEPS_PER_PATH
        run = eng.run_scan_async
        cfg = dict(workload=f"C2-HL: {P} device GBM OHLC paths x 390 per GPU, exits on bar high/low", bytes_per_step=12)
    elif name == "c4":
        P = args.paths or 1_000_000
        grid = [abi.make_params(contact_prox=0.10, stop_padding=0.10 + 0.05 * i, tp_padding=tp, cooldown_s=cd)
                for i in range(16) for tp in (0.15, 0.25, 0.35, 0.50) for cd in (0, 60, 300, 900)]
        eng.set_params(grid); eng.set_lr(lr_weights())
        eng.gen_gbm(F.HOST_SEED, rank * P, P, STEPS_PER_PATH, g)
        units = len(grid) * P * STEPS_PER_PATH
        run = eng.run_scan_async
        cfg = dict(workload=f"C4: {len(grid)} stop/target x cooldown combos x {P} resident GBM paths x 390 per GPU",
                   bytes_per_step=4.0 / len(grid), unit_note="combo-path-steps")
    elif name == "c5":
        P, days = args.paths or 1_250_000, 5
        rng = np.random.default_rng(5)
        lr5 = np.column_stack([rng.normal(0.2, 0.2, days), rng.normal(-0.5, 0.2, days), rng.normal(-0.45, 0.1, days),
                               rng.normal(0.35, 0.1, days), rng.normal(1.0, 0.1, days)])
        eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)); eng.set_lr(lr5)
        window = F.synthetic_window(40)
        T = days * STEPS_PER_PATH
        units = P * T

        def run():
            eng.gen_bootstrap(F.HOST_SEED, rank * P, P, T, window, 5)
            eng.run_scan_async()
        cfg = dict(workload=f"C5: block bootstrap (block 5) of a 40-bar window into {P} x {days}-day paths per GPU, "
                            "per-day LR weights, exits on high/low; generate (c,h,l) then scan", bytes_per_step=24)
    elif name == "c5f":
        P, days = args.paths or 1_250_000, 5
        rng = np.random.default_rng(5)
        lr5 = np.column_stack([rng.normal(0.2, 0.2, days), rng.normal(-0.5, 0.2, days), rng.normal(-0.45, 0.1, days),
                               rng.normal(0.35, 0.1, days), rng.normal(1.0, 0.1, days)])
        eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)); eng.set_lr(lr5)
        window = F.synthetic_window(40)
        T = days * STEPS_PER_PATH
        units = P * T
        run = lambda: eng.run_fused_bootstrap_async(F.HOST_SEED, rank * P, P, T, window, 5)
        cfg = dict(workload=f"C5 fused: block bootstrap (block 5) of a 40-bar window into {P} x {days}-day paths per GPU, "
                            "per-day LR weights, exits on high/low; generated and scanned in registers", bytes_per_step=0)
    else:
        raise SystemExit(f"unknown workload {name}")
    scratch = [None]

    def step():
        run()
        if world > 1:
            scratch[0] = D.allgather_merge_(eng, local_rank, scratch=scratch[0])
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    res = eng.fetch_agg()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank); sampler.start()
    l0 = eng.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clocks = sampler.stop()
    launches = eng.launch_count() - l0
    if rank == 0:
        peak, src = measured_peak_hbm()
        ach = cfg["bytes_per_step"] * units / (ms * 1e-3) / 1e9
        print(json.dumps({
            "metric": METRIC, "value": units * world / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 bars", "data": "synthetic", "config": cfg,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                         "note": "whole step (all kernels) / algorithmic bytes", "peak_source": src},
            "roofline_fp32": fp32_roofline(eng, units / (ms * 1e-3), 24 if name == "c3" else 12),
            "clocks": clocks, "gpu_launches": launches,
            "result_check": {"n_paths": res[0].n_paths, "n_trades": res[0].n_trades, "hit_rate": res[0].hit_rate,
                             "combos": len(res)}}), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def run_jitter_bench(args, rank, local_rank, world):
    """SURVEY 8(d) jitter-MC mode: the reference's own Monte Carlo (`simulate_monte_carlo`, M:3353-3538) — candidates
    found on the host over one 390-bar day, trials x candidates forward walks on the device (Philox draws).  Unit:
    candidate-trials/s (one jittered walk to its stop, target or the end of the day); the walks end at the first hit,
    so bars-visited is data dependent and only its upper bound (n - candidate bar - 1 per walk) is printed beside it.
    The reference spends ~6 us per candidate-trial in pure Python."""
    import torch
    from qmmx_monolithic_monte_carlo_b200 import engine as E, montecarlo as MC
    torch.cuda.set_device(local_rank)
    eng = E.MCEngine(local_rank, args.lib)
    w = F.synthetic_window(390, seed=11)
    c, h, l = (w[:, 3].astype(np.float64), w[:, 1].astype(np.float64), w[:, 2].astype(np.float64))
    rows, price, _ = E.sort_levels(F.levels_30())
    cands = MC.find_candidates(c, price, 0.10, 0.35, 0.25, 2)
    trials = args.paths or (1 << 22)
    sig = (0.02, 0.01, 0.0, 0.0)
    units = trials * sum(len(c) - cd["i"] - 1 for cd in cands)
    run = lambda: eng.run_jitter(h, l, cands, 0.35, 0.25, sig, trials, seed=7 + rank)
    for _ in range(args.warmup):
        run()
    sampler = ClockSampler(local_rank); sampler.start()
    l0 = eng.launch_count()
    ks, t0 = [], time.perf_counter()
    for _ in range(args.steps):
        out = run()
        ks.append(eng.last_timing_ms())
    wall = (time.perf_counter() - t0) / args.steps
    clocks = sampler.stop()
    launches = eng.launch_count() - l0
    k_ms = float(np.mean(ks))
    if rank == 0:
        s = MC.summarize(out["totals"], out["drawdowns"], out["wins"], out["losses"], out["opens"], len(cands))
        print(json.dumps({
            "metric": "jitter Monte Carlo candidate-trials/sec", "value": trials * len(cands) / (k_ms * 1e-3), "unit": "candidate-trials/s",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": k_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"jitter MC: {trials} trials x {len(cands)} candidates over one 390-bar day, 30 levels",
                       "timing": "device time of jitter_kernel (CUDA events on the ctx stream)",
                       "candidate_trial_steps_upper_bound_per_s": units / (k_ms * 1e-3)},
            "roofline": {"bound": "hbm", "achieved": 0.0, "peak": measured_peak_hbm()[0], "unit": "GB/s", "frac": 0.0, "traffic": None,
                         "note": "high/low live in shared memory: no HBM bytes per step (SURVEY 8d); outputs are 32 B per trial"},
            "e2e": {"value": trials * len(cands) / wall, "unit": "candidate-trials/s", "h2d_bytes_per_step": int(h.nbytes + l.nbytes),
                    "d2h_bytes_per_step": int(trials * 28), "ms_per_step": wall * 1e3,
                    "api": "MCEngine.run_jitter -> qmmx_run_jitter (host arrays in, per-trial totals out)"},
            "clocks": clocks, "gpu_launches": launches,
            "result_check": {"candidates": len(cands), "mean_R": s.mean, "var05": s.var05}}), flush=True)
    eng.close()


def _timed(eng, stream, world, local_rank, torch, dist, D, run, merge=True):
    """One launch sequence, device-timed on `stream`, one merge over the ranks at the end, max over ranks."""
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    run()
    if world > 1 and merge:
        D.allgather_merge_(eng, local_rank)
    e1.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sub_lines(args, eng, stream, rank, local_rank, world, torch, dist, D):
    """BASELINE.json configs[2..4] and the reference's own Monte Carlo, each run once (after a warm-up launch) outside
    the headline's timed region.  c3 is STRONG-scaled as BASELINE.json states it: `--c3-paths` (100M) global path ids
    split by distributed.shard_range, so every integer aggregate -- and `agg_digest` -- must be the same at N = 1, 2, 4, 8."""
    from oracle import c_oracle as CO
    from qmmx_monolithic_monte_carlo_b200 import engine as E, montecarlo as MC
    out = {}
    rows, price, solid = E.sort_levels(F.levels_30())
    fp32_peak = eng.measure_fp32_peak()
    # ---- c3: the target config -----------------------------------------------------------------------------------
    g = abi.make_gbm()
    params = engine_params()
    eng.set_levels(F.levels_30()); eng.set_params(params); eng.set_lr(lr_weights())
    total = args.c3_paths
    lo, hi = D.shard_range(total, rank, world)
    eng.run_fused_gbm_async(F.HOST_SEED, lo, min(hi - lo, 1 << 20), STEPS_PER_PATH, g)          # warm-up launch
    l0 = eng.launch_count()
    ms = _timed(eng, stream, world, local_rank, torch, dist, D,
                lambda: eng.run_fused_gbm_async(F.HOST_SEED, lo, hi - lo, STEPS_PER_PATH, g))
    res = eng.fetch_agg()
    assert res[0].n_paths == total and res[0].n_steps == total * STEPS_PER_PATH, (res[0].n_paths, total)
    units = total * STEPS_PER_PATH
    c3 = {"workload": f"C3: {total} Philox GBM paths x 390, strong-scaled over {world} GPU(s) by global path id, fused "
                      "generate+scan (no bar traffic), one all-gather + merge at the end",
          "value": units / (ms * 1e-3), "unit": UNIT, "ms": ms, "paths_total": total, "scaling": "strong",
          "gpu_launches": int(eng.launch_count() - l0), "agg_digest": agg_digest(res),
          "n_trades": res[0].n_trades, "n_wins": res[0].n_wins, "sum_pnl": res[0].sum_pnl,
          "roofline_fp32": {"achieved": 24 * units / (ms * 1e-3) / 1e12 / world, "peak": fp32_peak, "unit": "TFLOP/s per GPU",
                            "frac": 24 * units / (ms * 1e-3) / 1e12 / world / fp32_peak, "flop_per_path_step": 24}}
    if rank == 0:     # the fused kernel against the C oracle on the bars the materialising generator writes for the same ids
        S = 65_536
        fused = eng.run_fused_gbm(F.HOST_SEED, 0, S, STEPS_PER_PATH, g)[0].agg
        eng.gen_gbm(F.HOST_SEED, 0, S, STEPS_PER_PATH, g)
        bars = eng.download_bars(S, STEPS_PER_PATH)[0]
        o = CO.run_paths(bars, price, solid, params, lr=lr_weights(), n_threads=os.cpu_count() or 1)["agg"]
        c3["oracle_check"] = {"paths": S, "what": "fused kernel == C oracle over qmmx_gen_gbm bars of path ids [0, 65536)",
                              "equal": bool(int_aggs_equal(fused, o))}
        assert c3["oracle_check"]["equal"], "C3: fused kernel aggregates differ from the oracle"
    out["c3"] = c3
    # ---- c2 with the OnlinePolicy gate ON (the reference's own default, M:3090; synthetic weights) -----
    Pp = args.paths or 1_000_000
    prng = np.random.default_rng(3)
    eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, disable_policy_gate=False))
    eng.set_policy(prng.normal(0.3, 0.6, (1, 3, 7)))
    eng.gen_gbm(F.HOST_SEED, rank * Pp, Pp, STEPS_PER_PATH, g)
    eng.run_scan_async()
    ms = _timed(eng, stream, world, local_rank, torch, dist, D, eng.run_scan_async)
    res = eng.fetch_agg()
    out["c2_policy_gate"] = {"workload": f"C2 with DISABLE_POLICY_GATE=0: {Pp} resident GBM paths x 390 per GPU, float64 OnlinePolicy gate "
                                         "(Neumaier dot, M:3046-3109) after the LR filter",
                             "value": Pp * world * STEPS_PER_PATH / (ms * 1e-3), "unit": UNIT, "ms": ms, "scaling": "weak",
                             "n_trades": res[0].n_trades, "online_policy_skips": res[0].entry_hist["ONLINE_POLICY"]}
    eng.set_policy(None)
    # ---- c2hl: C2 with OHLC bars and exits on the bar's high / low (12 B per path-step; lockstep lean kernel) ------
    Ph = args.paths or 1_000_000
    eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True))
    eng.gen_gbm(F.HOST_SEED, rank * Ph, Ph, STEPS_PER_PATH, abi.make_gbm(sub_steps=4))
    eng.run_scan_async()
    hl_ms = [_timed(eng, stream, world, local_rank, torch, dist, D, eng.run_scan_async) for _ in range(5)]
    ms = float(np.median(hl_ms))
    res = eng.fetch_agg()
    peak, peak_src = measured_peak_hbm()
    gbs = 12.0 * Ph * STEPS_PER_PATH / (ms * 1e-3) / 1e9
    out["c2hl"] = {"workload": f"C2-HL: {Ph} device GBM OHLC paths x 390 per GPU (4 sub-steps per bar), exits on bar high/low, "
                               "stop before target (M:3621-3636); bars 4.68 GB per GPU, larger than L2",
                   "value": Ph * world * STEPS_PER_PATH / (ms * 1e-3), "unit": UNIT, "ms": ms, "scaling": "weak",
                   "n_trades": res[0].n_trades,
                   "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                                "algorithmic_bytes_per_launch": 12.0 * Ph * STEPS_PER_PATH, "peak_source": peak_src,
                                "kernel": "lean_scan_kernel<0, 1, 0, 0, 0, 0> (lockstep; the tile-event form was measured and loses, DESIGN.md section 8)",
                                "timing": "median of 5 launches, CUDA events on the launching stream"}}
    # ---- c4: 256-combo sweep over resident bars (weak: 1M paths per GPU) ----------------------------------------
    P4 = args.paths or 1_000_000
    grid = [abi.make_params(contact_prox=0.10, stop_padding=0.10 + 0.05 * i, tp_padding=tp, cooldown_s=cd)
            for i in range(16) for tp in (0.15, 0.25, 0.35, 0.50) for cd in (0, 60, 300, 900)]
    eng.set_params(grid)
    eng.gen_gbm(F.HOST_SEED, rank * P4, P4, STEPS_PER_PATH, g)
    eng.run_scan_async()
    l0 = eng.launch_count()
    ms = _timed(eng, stream, world, local_rank, torch, dist, D, eng.run_scan_async)
    res = eng.fetch_agg()
    units = len(grid) * P4 * world * STEPS_PER_PATH
    out["c4"] = {"workload": f"C4: {len(grid)} stop/target x cooldown combos x {P4} resident GBM paths x 390 per GPU (weak)",
                 "value": units / (ms * 1e-3), "unit": "combo-path-steps/s", "ms": ms, "scaling": "weak",
                 "gpu_launches": int(eng.launch_count() - l0), "agg_digest": agg_digest(res), "combos": len(res),
                 "n_trades_combo0": res[0].n_trades}
    # ---- c5f: fused block bootstrap of a 40-bar window into 5-day paths, per-day LR tables (weak: 1.25M per GPU) ---
    P5, days = (args.paths or 1_250_000), 5
    rng = np.random.default_rng(5)
    lr5 = np.column_stack([rng.normal(0.2, 0.2, days), rng.normal(-0.5, 0.2, days), rng.normal(-0.45, 0.1, days),
                           rng.normal(0.35, 0.1, days), rng.normal(1.0, 0.1, days)])
    eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)); eng.set_lr(lr5)
    window = F.synthetic_window(40)
    T5 = days * STEPS_PER_PATH
    eng.run_fused_bootstrap_async(F.HOST_SEED, rank * P5, min(P5, 1 << 17), T5, window, 5)
    l0 = eng.launch_count()
    ms = _timed(eng, stream, world, local_rank, torch, dist, D,
                lambda: eng.run_fused_bootstrap_async(F.HOST_SEED, rank * P5, P5, T5, window, 5))
    res = eng.fetch_agg()
    units = P5 * world * T5
    out["c5f"] = {"workload": f"C5 fused: block bootstrap (block 5) of a 40-bar window into {P5} x {days}-day paths per GPU (weak), "
                              "per-day LR weights, exits on high/low",
                  "value": units / (ms * 1e-3), "unit": UNIT, "ms": ms, "scaling": "weak",
                  "gpu_launches": int(eng.launch_count() - l0), "agg_digest": agg_digest(res), "n_trades": res[0].n_trades}
    # ---- mc: the reference's own jitter Monte Carlo (rank 0's GPU) -----------------------------------------------
    if rank == 0:
        w = F.synthetic_window(390, seed=11)
        c, h, l = (w[:, 3].astype(np.float64), w[:, 1].astype(np.float64), w[:, 2].astype(np.float64))
        gates = MC.EntryGates(rows, price, solid, contact_prox=0.10, stop_padding=0.35, lr=lr_weights()[0])
        cands = MC.find_candidates(c, price, 0.10, 0.35, 0.25, 2, gates=gates, gate_order="mc")     # with_gates=True
        trials = 1 << 22
        eng.run_jitter(h, l, cands, 0.35, 0.25, (0.02, 0.01, 0.0, 0.0), 1 << 16, seed=7)
        l0 = eng.launch_count()
        o = eng.run_jitter(h, l, cands, 0.35, 0.25, (0.02, 0.01, 0.0, 0.0), trials, seed=7)
        k_ms = eng.last_timing_ms()
        sm = MC.summarize(o["totals"], o["drawdowns"], o["wins"], o["losses"], o["opens"], len(cands))
        ref_t = python_reference_timing()
        out["mc"] = {"workload": f"jitter MC (simulate_monte_carlo, with_gates=True): {trials} trials x {len(cands)} candidates over one 390-bar day",
                     "value": trials * len(cands) / (k_ms * 1e-3), "unit": "candidate-trials/s", "ms": k_ms, "n_gpus": 1,
                     "gpu_launches": int(eng.launch_count() - l0), "mean_R": sm.mean, "var05": sm.var05,
                     "reference_python": (ref_t or {}).get("simulate_monte_carlo")}
    return out


def python_reference_timing():
    try:
        with open(os.path.join(ROOT, "profiles", "python_reference_timing.json")) as f:
            return json.load(f)
    except Exception:
        return None


def _nb_worker(job):
    """one process of the all-cores NumPy batch baseline: its own GBM paths (the headline recipe, its own seed)"""
    seed, n, levels, lr_row = job
    from oracle import headless as H, numpy_batch as NB
    lv = H.sort_levels(levels)
    P = H.Params(contact_prox=0.10, stop_padding=0.35, tp_padding=0.25, q_min_prob=0.60, cooldown_s=120)
    m = H.LRModel(list(lr_row[:4]), float(lr_row[4])) if lr_row is not None else None
    bars = F.host_gbm_closes(n, STEPS_PER_PATH, seed=seed)
    t0 = time.perf_counter()
    NB.run_batch(bars, lv, P, m)
    return time.perf_counter() - t0


def python_baselines(close, levels, lr):
    """The Python/NumPy CPU baselines SURVEY 8(d) names, on bounded samples of the headline bars (rank 0, host cores):
    the NumPy batch headless engine (oracle/numpy_batch.py), the scalar headless function (oracle/headless.run_path), and
    -- measured in the build container, because the reference does not travel to the GPU box -- the unmodified
    reference's own evaluate_entry."""
    from oracle import headless as H, numpy_batch as NB
    lv = H.sort_levels(levels)
    P = H.Params(contact_prox=0.10, stop_padding=0.35, tp_padding=0.25, q_min_prob=0.60, cooldown_s=120)
    m = H.LRModel(list(lr[0][:4]), float(lr[0][4])) if lr is not None else None
    nb_paths = min(close.shape[1], 20_000)
    sample = np.ascontiguousarray(close[:, :nb_paths])
    t0 = time.perf_counter(); NB.run_batch(sample, lv, P, m); dt_nb = time.perf_counter() - t0
    sc_paths = min(close.shape[1], 1000)
    t0 = time.perf_counter()
    for j in range(sc_paths):
        H.run_path(close[:, j], lv, P, lr=[m] if m else None)
    dt_sc = time.perf_counter() - t0
    out = {"numpy_batch": {"value": nb_paths * STEPS_PER_PATH / dt_nb, "unit": UNIT, "cores": 1,
                           "sample": f"first {nb_paths} paths, oracle/numpy_batch.py (vectorised over paths, Python loop over steps)"},
           "headless_scalar": {"value": sc_paths * STEPS_PER_PATH / dt_sc, "unit": UNIT, "cores": 1,
                               "sample": f"first {sc_paths} paths, oracle/headless.run_path (pure Python)"},
           "host_cpu": cpu_model(), "host_logical_cores": os.cpu_count()}
    try:        # the same NumPy batch engine on every core: spawned workers, no CUDA in them
        import multiprocessing as mp
        cores = os.cpu_count() or 1
        per = 10_000
        jobs = [(1000 + k, per, levels, None if lr is None else [float(x) for x in lr[0]]) for k in range(cores)]
        with mp.get_context("spawn").Pool(cores) as pool:
            pool.map_async(_nb_worker, [(1, 200, levels, jobs[0][3])] * cores).get(timeout=90)     # import + warm-up
            t0 = time.perf_counter()
            pool.map_async(_nb_worker, jobs, chunksize=1).get(timeout=120)      # a stuck worker must not stall the line
            dt_all = time.perf_counter() - t0
        out["numpy_batch_all_cores"] = {"value": cores * per * STEPS_PER_PATH / dt_all, "unit": UNIT, "cores": cores,
                                        "sample": f"{cores} processes x {per} GBM paths of the headline recipe (own seeds), oracle/numpy_batch.py"}
    except Exception as e:
        out["numpy_batch_all_cores"] = {"error": str(e)[:120]}
    ref = python_reference_timing()
    if ref:
        out["reference_evaluate_entry"] = {"value": ref["evaluate_entry"]["path_steps_per_s"], "unit": UNIT, "cores": 1,
                                           "sample": f"{ref['evaluate_entry']['paths']} paths, " + ref["evaluate_entry"]["what"],
                                           "measured_on": ref["measured_on"] + " (profiles/python_reference_timing.json)"}
    return out


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--paths", type=int, default=0, help="override paths per GPU (debug)")
    ap.add_argument("--c3-paths", type=int, default=100_000_000, help="global path count of the strong-scaled c3 sub-line")
    ap.add_argument("--no-sub", action="store_true", help="skip the c3 / c4 / c5f / mc sub-lines")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--lib", default=None, help="alternate libqmmx_mc build (kernel tuning variants)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "ours" and args.workload == "mc":
        run_jitter_bench(args, rank, local_rank, world)
        return
    if args.impl == "ours" and args.workload in ("c3", "c4", "c5", "c5f", "c2hl"):
        run_extra(args, args.workload, rank, local_rank, world)
        return
    wl = workload(args.workload)
    if args.paths:
        wl["paths"] = args.paths

    if args.impl == "reference":
        run_reference(args, wl, rank)
        return

    import torch
    import torch.distributed as dist
    from qmmx_monolithic_monte_carlo_b200 import engine as E

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    numa = bind_numa(local_rank, world)      # before the host bars exist: first touch puts them on the rank's node
    paths = wl["paths"]
    rows, price, solid = E.sort_levels(wl["levels"])
    params = engine_params()
    lr = lr_weights() if wl["lr"] else None
    close = F.host_gbm_closes(paths, STEPS_PER_PATH, seed=F.HOST_SEED + rank)   # weak scaling: own paths per rank

    eng = E.MCEngine(local_rank, args.lib)
    stream = torch.cuda.Stream()           # non-default: its handle is non-NULL, so the ctx can adopt it
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)
    eng.set_levels(wl["levels"])
    eng.set_params(params)
    eng.set_lr(lr)
    eng.upload_bars(close)

    from qmmx_monolithic_monte_carlo_b200 import distributed as D

    scratch = [None]

    def step():
        eng.run_scan_async()
        if world > 1:   # a step's result is the whole job's aggregates: one NCCL all-gather of the 2.3 KB vector + the
            scratch[0] = D.allgather_merge_(eng, local_rank, scratch=scratch[0])      # library's merge kernel

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    res = eng.fetch_agg()[0]
    expect_paths = paths * world
    assert res.n_paths == expect_paths and res.n_steps == expect_paths * STEPS_PER_PATH, (res.n_paths, expect_paths)

    sampler = ClockSampler(local_rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    launches0 = eng.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total_ms = e0.elapsed_time(e1)
    launches = eng.launch_count() - launches0
    clocks = sampler.stop()
    # per-launch duration of the dominant kernel: the library's own event pair on the same stream
    for _ in range(min(args.steps, 10)):
        eng.run_scan_async()
        torch.cuda.synchronize()
        kernel_ms.append(eng.last_timing_ms())
    t = torch.tensor([total_ms], dtype=torch.float64, device=f"cuda:{local_rank}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = expect_paths * STEPS_PER_PATH / (ms_per_step * 1e-3)

    # ---- end to end through the public API with HOST buffers (pinned), H2D inside the timed region ----
    e2e = None
    if not args.no_e2e:
        eng.pin(close)
        try:
            n_e2e = max(2, min(args.steps, 5))
            eng.run_scan_host(close)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                r = eng.run_scan_host(close)[0]           # returns the aggregates on the host (D2H inside)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / n_e2e
            tt = torch.tensor([dt], dtype=torch.float64, device=f"cuda:{local_rank}")
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
            assert r.n_paths == paths
            e2e = {"value": expect_paths * STEPS_PER_PATH / dt, "unit": UNIT, "h2d_bytes_per_step": int(close.nbytes),
                   "d2h_bytes_per_step": int(abi.AGG_SLOTS * 8), "ms_per_step": dt * 1e3,
                   "api": "MCEngine.run_scan_host -> qmmx_run_scan_host (pinned host bars, chunked H2D overlapped with the scan)"}
            # ceiling of the host side: the same pinned bars as ONE plain copy per rank (qmmx_upload_bars: cudaMemcpy2DAsync
            # with pitch == width), all ranks at once.  e2e cannot beat it; the gap is what the pipeline leaves on the table.
            eng.upload_bars(close)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            for _ in range(2):
                eng.upload_bars(close)
            pt = (time.perf_counter() - t0) / 2
            tp = torch.tensor([pt], dtype=torch.float64, device=f"cuda:{local_rank}")
            if world > 1:
                dist.all_reduce(tp, op=dist.ReduceOp.MAX)
            pt = float(tp.item())
            e2e["h2d_probe"] = {"gb_per_s_per_rank": close.nbytes / pt / 1e9, "gb_per_s_aggregate": close.nbytes * world / pt / 1e9,
                                "ms": pt * 1e3, "what": "plain pinned H2D of the same bars, all ranks concurrently (slowest rank)",
                                "e2e_over_probe": pt / dt}
            e2e["numa"] = numa
        finally:
            eng.unpin(close)

    if rank == 0:
        peak, peak_src = measured_peak_hbm()
        k_ms = float(np.mean(kernel_ms))
        cap = shipped_capture()
        # the binding resource against a MEASURED ceiling: warp instructions per second of the scan (instruction counts of
        # the committed capture of the same kernel on the same workload / the launch time measured in this run) next to
        # what LOP3-only and LOP3 + FFMA probe kernels reach on this device in this run
        issue = None
        try:
            alu_probe, mixed_probe = eng.measure_issue_peak()
            if cap and paths == cap.get("paths") and cap.get("inst_executed"):
                tot = cap["inst_executed"] / (k_ms * 1e-3) / 1e9
                alu = cap["alu_inst_executed"] / (k_ms * 1e-3) / 1e9
                issue = {"unit": "1e9 warp-instructions/s, whole device",
                         "alu_pipe": {"achieved": alu, "probe": alu_probe, "frac": alu / alu_probe},
                         "all_pipes": {"achieved": tot, "probe_alu_plus_fma": mixed_probe, "frac": tot / mixed_probe},
                         "probe_source": "qmmx_measure_issue_peak: LOP3-only and LOP3 + FFMA kernels, this run"}
            else:
                issue = {"unit": "1e9 warp-instructions/s, whole device", "probe_alu": alu_probe, "probe_alu_plus_fma": mixed_probe}
        except Exception as e:      # an older library without the probe
            issue = {"error": str(e)}
        alg_bytes = 4.0 * paths * STEPS_PER_PATH
        achieved = alg_bytes / (k_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 bars; f64-exact decisions via fp32 thresholds; f64 PnL sums",
            "data": "synthetic",
            "config": headline_config(wl, world, len(rows)),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (cap or {}).get("dram_bytes") if paths == 1_000_000 else None,
                         "traffic_source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one launch (%s)" %
                                           (cap or {}).get("summary", "no committed capture"),
                         "kernel": (cap or {}).get("kernel", "tile_scan_tma_kernel<MULTIDAY=0,SLOW=0>"), "kernel_ms": k_ms,
                         "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                         "binding_limit": {"what": "instruction issue (and the ALU pipe), not HBM",
                                           "issue_active_pct": (cap or {}).get("issue_active_pct"),
                                           "warp_instructions_per_warp_step": (cap or {}).get("warp_instructions_per_warp_step"),
                                           "active_lanes_per_instruction": (cap or {}).get("active_lanes_per_instruction"),
                                           "alu_pipe_pct": (cap or {}).get("alu_pipe_pct"),
                                           "measured_issue_rates": issue,
                                           "source": "committed ncu capture of the shipped build: %s at git %s (not re-measured per run)" %
                                                     ((cap or {}).get("summary"), (cap or {}).get("git"))}},
            "roofline_fp32": fp32_roofline(eng, paths * STEPS_PER_PATH / (k_ms * 1e-3), 12),
            "clocks": clocks, "gpu_launches": int(launches),
            "result_check": {"n_trades": res.n_trades, "n_wins": res.n_wins, "sum_pnl": res.sum_pnl},
        }
        if e2e:
            line["e2e"] = e2e
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            sample = min(paths, 262_144)
            sample_bars = np.ascontiguousarray(close[:, :sample])      # contiguous BEFORE the timer starts
            agg, dt = cpu_port_run(sample_bars, price, solid, params, lr, threads)
            line["cpu_baseline"] = {"value": sample * STEPS_PER_PATH / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"first {sample} paths x {STEPS_PER_PATH} steps of the same bars, "
                                              "oracle/qmmx_oracle.c with pthreads", "host_cpu": cpu_model()}
            # the GPU scan of the same sample must give the oracle's aggregates (every counter, both histograms, PnL sum)
            eng.upload_bars(sample_bars)
            g_agg = eng.run_scan()[0].agg
            line["oracle_check"] = {"paths": sample, "what": "GPU scan == C oracle on the first 262144 paths of the timed bars",
                                    "equal": bool(int_aggs_equal(g_agg, agg))}
            assert line["oracle_check"]["equal"], "C2: GPU aggregates differ from the oracle on the sample"
            line["python_baselines"] = python_baselines(close, wl["levels"], lr)
    subs = None
    if not args.no_sub and args.workload == "c2":
        subs = sub_lines(args, eng, stream, rank, local_rank, world, torch, dist, D)
    if rank == 0:
        if subs:
            line.update(subs)
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
