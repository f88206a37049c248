"""Times the UNMODIFIED reference's own Python on this container's host cores (needs /root/reference, so it cannot
run on the GPU box): the stub-imported `MonolithicEngine.evaluate_entry` driven tick by tick over C2-style paths
(M:1492, with its SQLite policy_events logging) and `QMMXApp.simulate_monte_carlo` (M:3353) on a 390-bar day.
Writes profiles/python_reference_timing.json, which bench.py quotes beside the numbers it measures on the GPU box.

    python profiles/time_python_baselines.py
"""
import json
import os
import platform
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import make_golden as MG, ref_harness as rh  # noqa: E402
from qmmx_monolithic_monte_carlo_b200 import fixtures as F  # noqa: E402


def cpu_model():
    try:
        for line in subprocess.run(["lscpu"], capture_output=True, text=True).stdout.splitlines():
            if line.startswith("Model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return platform.processor()


def main():
    paths, steps = 100, 390
    close = F.host_gbm_closes(paths, steps)
    levels = F.levels_30()
    settings = {"CONTACT_PROX": "0.1", "Q_SIGNAL_COOLDOWN": "120", "ENABLE_VETO": "0"}
    model = rh.make_sklearn_model([0.2, -0.5, -0.45, 0.35], 1.0)
    rh.drive_reference_path(close[:, 0], levels, settings, lr_model=model)          # import + warm-up
    t0 = time.perf_counter()
    for j in range(paths):
        rh.drive_reference_path(close[:, j], levels, settings, lr_model=model)
    dt_eval = time.perf_counter() - t0
    bars = MG.synth_ohlc(390, 21)
    dict_bars = [dict(t=60000 * i, o=bars["o"][i], h=bars["h"][i], l=bars["l"][i], c=bars["c"][i]) for i in range(390)]
    mc = {}
    for wg in (False, True):
        t0 = time.perf_counter()
        msg = rh.run_reference_monte_carlo(dict_bars, levels, {"CONTACT_PROX": "0.1", "ENABLE_VETO": "0"}, n=390, touch_limit=2,
                                           trials=500, with_gates=wg, seed=7)
        dt = time.perf_counter() - t0
        cands = int(msg[2].split("Candidates: ")[1].split(" ")[0])
        mc["with_gates_%s" % wg] = {"seconds": dt, "candidates": cands, "trials": 500,
                                    "candidate_trials_per_s": cands * 500 / dt}
    out = {
        "measured_on": "build container (no GPU), %s, %d logical cores, single process" % (cpu_model(), os.cpu_count()),
        "python": platform.python_version(),
        "evaluate_entry": {"what": "stub-imported reference MonolithicEngine.evaluate_entry (M:1492) + exit block, one engine "
                                   "per path on sqlite :memory:, sklearn LogisticRegression model set, 30 levels",
                           "paths": paths, "steps": steps, "seconds": dt_eval, "path_steps_per_s": paths * steps / dt_eval, "cores": 1},
        "simulate_monte_carlo": {"what": "reference QMMXApp.simulate_monte_carlo (M:3353), n=390, trials=500, touch_limit=2, 30 levels", **mc},
    }
    with open(os.path.join(ROOT, "profiles", "python_reference_timing.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
