"""Design-time simulation (CPU, oracle traces) of warp schedules for the engine scan.

Not a product path and not a benchmark: it replays the per-step reason codes and trade records the C oracle
produces for the C2 workload and counts MODELLED warp-instructions per warp-step for candidate kernel
structures, so that the structure is chosen before a kernel is written.

    python profiles/sim_event_schedule.py [paths]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import c_oracle as CO  # noqa: E402
from qmmx_monolithic_monte_carlo_b200 import abi, engine as E, fixtures as F  # noqa: E402

# modelled instruction costs (warp-instructions per executed block)
P1 = 8        # phase 1 per step: load, stage, cell index, cell byte, mask bit (all lanes active, unrolled)
C = 44        # contact block (latch, touch byte, gate record, entry arm)
S = 9         # one exit-test step of a lane in a position
X = 24        # exit block
V = 8         # vote / loop overhead per round


def traces(paths, hl=False):
    rows, price, solid = E.sort_levels(F.levels_30())
    params = abi.make_params(contact_prox=0.10, stop_padding=0.35, tp_padding=0.25, q_min_prob=0.60, cooldown_s=120)
    lr = np.array([[0.2, -0.5, -0.45, 0.35, 1.0]])
    close = F.host_gbm_closes(paths, 390)
    o = CO.run_paths(close, price, solid, params, lr=lr, want_codes=True, max_trades=64, n_threads=8)
    codes = o["codes"]                                   # [steps][paths]
    code = {n: i for i, n in enumerate(abi.ENTRY_CODES)}
    quiet = [code[n] for n in ("TOO_FAR", "IN_POSITION", "COOLDOWN", "DIR_UNKNOWN", "NOLEVELS")]
    contact = ~np.isin(codes, quiet)                     # evaluate_entry went past the nearest-level gate
    inpos = codes == code["IN_POSITION"]                     # exit-tested steps (position open at the start of the step)
    exits = np.zeros_like(contact)
    tr, cnt = o["trades"], o["counts"]
    for p in range(paths):
        for t in range(min(cnt[p], 64)):
            xb = tr[p, t]["exit_bar"]
            if xb >= 0:
                exits[xb, p] = True
    return contact, inpos, exits


def lockstep(contact, inpos, exits):
    """today's kernel: a warp-step costs the base plus every branch some lane takes"""
    steps, paths = contact.shape
    w = paths // 32
    c = contact[:, : w * 32].reshape(steps, w, 32).any(axis=2)
    x = exits[:, : w * 32].reshape(steps, w, 32).any(axis=2)
    base = 17
    return base + C * c.mean() + X * x.mean()


def tile_rounds(contact, inpos, exits, R, window=None, thr=0):
    """Phase 1 in lockstep per tile of R steps (contact masks), phase 2 in rounds: every round each lane that is in a
    position tests ONE step (S), lanes with a pending exit run X together, flat lanes jump to their next eligible
    contact and run C together.  `window` (> R): a lane may lag up to window - R steps behind the tile being loaded
    (sliding ring); None = barrier at every tile end.  thr = minimum lanes for the C block unless nothing else can run."""
    steps, paths = contact.shape
    w = paths // 32
    total = 0.0
    # per-lane event lists in step order: (step, kind) kind 0 = exit-test step, 1 = contact, 2 = exit (after its test)
    ev = []
    for p in range(w * 32):
        l = []
        for s in np.flatnonzero(contact[:, p] | inpos[:, p]):
            if inpos[s, p]:
                l.append((s, 0))
                if exits[s, p]:
                    l.append((s, 2))
            else:
                l.append((s, 1))
        ev.append(l)
    for wi in range(w):
        lanes = ev[wi * 32:(wi + 1) * 32]
        ptr = [0] * 32
        lag = 0 if window is None else window - R
        cost = 0.0
        for t0 in range(0, steps, R):
            t1 = min(steps, t0 + R)
            cost += P1 * (t1 - t0)
            # a lane must finish everything before `must` now; may work up to t1
            must = t1 if window is None else max(0, t1 - lag)
            if t1 == steps:
                must = steps
            while True:
                kinds = [lanes[k][ptr[k]][1] if ptr[k] < len(lanes[k]) and lanes[k][ptr[k]][0] < t1 else -1 for k in range(32)]
                behind = any(ptr[k] < len(lanes[k]) and lanes[k][ptr[k]][0] < must for k in range(32))
                if not behind:
                    break
                n0, n1, n2 = kinds.count(0), kinds.count(1), kinds.count(2)
                cost += V
                ran = False
                if n0:
                    cost += S; ran = True
                    for k in range(32):
                        if kinds[k] == 0: ptr[k] += 1
                if n2:
                    cost += X; ran = True
                    for k in range(32):
                        if kinds[k] == 2: ptr[k] += 1
                if n1 and (n1 >= thr or not ran):
                    cost += C
                    for k in range(32):
                        if kinds[k] == 1: ptr[k] += 1
        total += cost
    return total / (w * steps)


if __name__ == "__main__":
    paths = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    c, i, x = traces(paths)
    print("per path: eligible contacts %.1f, exit-test steps %.1f, exits %.1f" % (c.sum() / paths, i.sum() / paths, x.sum() / paths))
    print("lockstep (model)            %.1f instr / warp-step" % lockstep(c, i, x))
    for R, win, thr in ((32, None, 0), (64, None, 0), (32, 64, 0), (32, 96, 0), (32, 64, 8), (64, 128, 0), (32, 128, 0), (32, 128, 10)):
        print("tile R=%-3d window=%-5s thr=%-2d %.1f" % (R, win, thr, tile_rounds(c, i, x, R, win, thr)))


def tile_rounds_b(contact, inpos, exits, R, window=None, K=4, P1_=P1):
    """Round structure B: one round = [C for flat lanes at their next eligible contact] then [up to K exit-test
    iterations for the lanes in a position, including the ones that just entered] then [X for the lanes that hit]."""
    steps, paths = contact.shape
    w = paths // 32
    ev = []
    for p in range(w * 32):
        l = []
        for s in np.flatnonzero(contact[:, p] | inpos[:, p]):
            if inpos[s, p]:
                l.append((s, 0))
                if exits[s, p]:
                    l.append((s, 2))
            else:
                l.append((s, 1))
        ev.append(l)
    total, rounds, cact = 0.0, 0, 0
    for wi in range(w):
        lanes = ev[wi * 32:(wi + 1) * 32]
        ptr = [0] * 32
        lag = 0 if window is None else window - R
        cost = 0.0
        for t0 in range(0, steps, R):
            t1 = min(steps, t0 + R)
            cost += P1_ * (t1 - t0)
            must = t1 if window is None else max(0, t1 - lag)
            if t1 == steps:
                must = steps

            def nxt(k):
                return lanes[k][ptr[k]] if ptr[k] < len(lanes[k]) and lanes[k][ptr[k]][0] < t1 else (None, -1)
            while any(ptr[k] < len(lanes[k]) and lanes[k][ptr[k]][0] < must for k in range(32)):
                cost += V
                rounds += 1
                cl = [k for k in range(32) if nxt(k)[1] == 1]
                if cl:
                    cost += C; cact += len(cl)
                    for k in cl: ptr[k] += 1
                pend = set()
                for it in range(K):
                    sl = [k for k in range(32) if k not in pend and nxt(k)[1] == 0]
                    if not sl:
                        break
                    cost += S
                    for k in sl:
                        ptr[k] += 1
                        if nxt(k)[1] == 2:
                            pend.add(k)
                if pend:
                    cost += X
                    for k in pend: ptr[k] += 1
        total += cost
    return total / (w * steps), rounds / (w * steps), cact / max(1, rounds)


if __name__ == "__main__":
    for R, win, K in ((32, None, 2), (32, None, 4), (32, None, 8), (64, None, 4), (32, 64, 4), (32, 64, 8), (32, 128, 4)):
        r = tile_rounds_b(c, i, x, R, win, K)
        print("B: tile R=%-3d window=%-5s K=%d  %.1f instr/step, %.3f rounds/step, %.1f lanes in C" % ((R, win, K) + r))


def tile_rounds_c(contact, inpos, exits, R, window, KB=2, costs=(16, 75, 21, 19, 14.0, 7.0)):
    """Structure C: contact rounds are batched (loop A: only the contact block, lanes that entered a position wait),
    then the exit tests (loop B: up to KB iterations of two rows per lane, the exit block when a lane hits), and
    again while a lane is behind.  costs = (head, contact, exit test, exit, phase 1 per step, chunk overhead per step),
    taken from the SASS counters of the built kernel (profiles/)."""
    V_, C_, S_, X_, P1_, OV_ = costs
    steps, paths = contact.shape
    w = paths // 32
    ev = []
    for p in range(w * 32):
        l = []
        for s in np.flatnonzero(contact[:, p] | inpos[:, p]):
            if inpos[s, p]:
                l.append((s, 0))
                if exits[s, p]:
                    l.append((s, 2))
            else:
                l.append((s, 1))
        ev.append(l)
    total, nA, nB = 0.0, 0, 0
    for wi in range(w):
        lanes = ev[wi * 32:(wi + 1) * 32]
        ptr = [0] * 32
        lag = window - R
        cost = 0.0
        for t0 in range(0, steps, R):
            t1 = min(steps, t0 + R)
            cost += (P1_ + OV_) * (t1 - t0)
            must = steps if t1 == steps else max(0, t1 - lag)

            def nxt(k):
                return lanes[k][ptr[k]] if ptr[k] < len(lanes[k]) and lanes[k][ptr[k]][0] < t1 else (None, -1)
            while any(ptr[k] < len(lanes[k]) and lanes[k][ptr[k]][0] < must for k in range(32)):
                # loop A
                while any(nxt(k)[1] == 1 and nxt(k)[0] < must for k in range(32)):
                    cost += V_ + C_; nA += 1
                    for k in range(32):
                        if nxt(k)[1] == 1: ptr[k] += 1
                cost += V_
                # loop B
                for it in range(KB):
                    sl = [k for k in range(32) if nxt(k)[1] == 0]
                    if not sl:
                        break
                    cost += S_; nB += 1
                    hit = False
                    for k in sl:
                        for rr in range(2):               # two rows per iteration
                            if nxt(k)[1] == 0:
                                ptr[k] += 1
                                if nxt(k)[1] == 2:
                                    ptr[k] += 1; hit = True
                                    break
                    if hit:
                        cost += X_
        total += cost
    return total / (w * steps), nA / (w * steps), nB / (w * steps)


if __name__ == "__main__":
    print("structure B with measured block costs (head 16, contact 75, exit test 21, exit 19, phase 1 14+7): %.1f" %
          (lambda r: 21 + r[1] * (16 + 75 + 21 + 0.76 * 19))(tile_rounds_b(c, i, x, 8, 24, 2)))
    for KB in (1, 2, 3):
        print("C: chunk 8 window 24 KB=%d  %.1f instr/step, %.3f contact rounds/step, %.3f exit iterations/step" % ((KB,) + tile_rounds_c(c, i, x, 8, 24, KB)))
