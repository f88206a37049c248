"""One small invocation per kernel, for ncu captures (profiles/capture.sh).  Not a benchmark: sizes are chosen so
that one launch of the kernel of interest runs a few ms and `ncu --set full` replays stay short.

    python profiles/prof_case.py <case> [paths]

cases: scan (lean/tile scan, C2 close-only), scanhl (C2-HL), sweep (C4, 32 combos), fused (C3), boot (C5 fused),
genboot (C5 generate), gengbm, jitter, lastbars, regime, retrain, policy (scan with the OnlinePolicy gate on)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qmmx_monolithic_monte_carlo_b200 import abi, engine as E, fixtures as F, montecarlo as MC  # noqa: E402

LR = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])


def main():
    case = sys.argv[1]
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    g = abi.make_gbm()
    with E.MCEngine(0) as eng:
        eng.set_levels(F.levels_30())
        eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120))
        eng.set_lr(LR)
        reps = 3
        if case == "scan":
            eng.gen_gbm(F.HOST_SEED, 0, P or 1_000_000, 390, g)
            run = eng.run_scan_async
        elif case == "policy":
            rng = np.random.default_rng(3)
            eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, disable_policy_gate=False))
            eng.set_policy(rng.normal(0.3, 0.6, (1, 3, 7)))
            eng.gen_gbm(F.HOST_SEED, 0, P or 1_000_000, 390, g)
            run = eng.run_scan_async
        elif case == "scanhl":
            eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True))
            eng.gen_gbm(F.HOST_SEED, 0, P or 1_000_000, 390, abi.make_gbm(sub_steps=4))
            run = eng.run_scan_async
        elif case == "sweep":
            grid = [abi.make_params(contact_prox=0.10, stop_padding=0.10 + 0.05 * i, tp_padding=tp, cooldown_s=120)
                    for i in range(16) for tp in (0.15, 0.25)]
            eng.set_params(grid)
            eng.gen_gbm(F.HOST_SEED, 0, P or 1_000_000, 390, g)
            run = eng.run_scan_async
        elif case == "fused":
            run = lambda: eng.run_fused_gbm_async(F.HOST_SEED, 0, P or 2_000_000, 390, g)
        elif case in ("boot", "genboot"):
            days = 5
            rng = np.random.default_rng(5)
            lr5 = np.column_stack([rng.normal(0.2, 0.2, days), rng.normal(-0.5, 0.2, days), rng.normal(-0.45, 0.1, days),
                                   rng.normal(0.35, 0.1, days), rng.normal(1.0, 0.1, days)])
            eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, exit_on_hl=True)); eng.set_lr(lr5)
            w = F.synthetic_window(40)
            if case == "boot":
                run = lambda: eng.run_fused_bootstrap_async(F.HOST_SEED, 0, P or 500_000, days * 390, w, 5)
            else:
                run = lambda: eng.gen_bootstrap(F.HOST_SEED, 0, P or 500_000, days * 390, w, 5)
        elif case == "gengbm":
            run = lambda: eng.gen_gbm(F.HOST_SEED, 0, P or 1_000_000, 390, abi.make_gbm(sub_steps=4))
        elif case in ("jitter", "lastbars"):
            w = F.synthetic_window(390, seed=11)
            c, h, l = (w[:, 3].astype(np.float64), w[:, 1].astype(np.float64), w[:, 2].astype(np.float64))
            rows, price, _ = E.sort_levels(F.levels_30())
            cands = MC.find_candidates(c, price, 0.10, 0.35, 0.25, 2)
            if case == "jitter":
                run = lambda: eng.run_jitter(h, l, cands, 0.35, 0.25, (0.02, 0.01, 0.0, 0.0), P or (1 << 20), seed=7)
            else:
                run = lambda: eng.run_last_bars(h, l, cands * 200)
        elif case == "regime":
            n = P or 200_000
            cl, hi, lo, vol = _boot_with_volume(n)
            eng.upload_bars(cl)
            run = lambda: eng.upload_volume(vol)          # regime_kernel runs when the volume plane arrives
        elif case == "retrain":
            n = P or 200_000
            rng = np.random.default_rng(3)
            eng.set_params(abi.make_params(contact_prox=0.10, cooldown_s=120, disable_policy_gate=False, online_retrain=True))
            eng.set_policy(rng.normal(0.3, 0.6, (1, 3, 7)))
            eng.gen_gbm(F.HOST_SEED, 0, n, 390, g)
            run = eng.run_scan_async
        else:
            raise SystemExit("unknown case " + case)
        for _ in range(reps):
            run()
        try:
            ms = eng.last_timing_ms()      # synchronises on the ctx stream's end event
        except Exception:
            ms = float("nan")              # calls that record no event pair (upload_volume)
            eng.fetch_agg() if case not in ("regime",) else None
        print("case %s: last launch %.3f ms, launches %d" % (case, ms, eng.launch_count()))


def _boot_with_volume(n):
    from qmmx_monolithic_monte_carlo_b200 import fixtures as FX
    rng = np.random.default_rng(9)
    steps = 390
    close = FX.host_gbm_closes(n, steps, sigma=0.02)
    vol = rng.lognormal(10.0, 0.6, size=(steps, n)).astype(np.float32)
    return close, None, None, vol


if __name__ == "__main__":
    main()
