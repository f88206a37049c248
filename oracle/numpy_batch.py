"""TEST INFRASTRUCTURE ONLY — NumPy batch restatement of the headless engine.

The same per-tick semantics as `oracle/headless.py:run_path` (M:1492-1771 gates, M:2927-3014 tick order and exits,
M:1926 pnl), vectorised over paths with a Python loop over the steps: the "headless Python/NumPy reference" the
north star asks to be timed on the box's host cores beside the GPU numbers.  Scope = BASELINE configs 1 and 2:
close-only exits, one logistic model in AND mode (or none); blend, OnlinePolicy, veto, multi-day tables and
high/low exits stay with headless.py / qmmx_oracle.c.  Checked against the C oracle in tests/test_c_oracle.py.
"""
from __future__ import annotations

import numpy as np

from oracle import headless as H

E = H.E
X = H.X


def run_batch(close, levels, params: "H.Params", lr: "H.LRModel | None" = None):
    """close: [steps][paths] fp32 (evaluated in float64, as the reference does).  Returns dict(entry_hist[20],
    exit_hist[4], n_trades, n_wins, n_open, sum_pnl, sum_pnl2, path_pnl[paths])."""
    P = params
    if P.use_blend or not P.disable_policy_gate or P.enable_veto or P.exit_on_hl or P.online_retrain:
        raise NotImplementedError("numpy_batch covers the AND-mode close-only engine (BASELINE configs 1-2)")
    close = np.asarray(close)
    steps, n = close.shape
    lp = np.array([lv.price for lv in levels], dtype=np.float64)
    solid = np.array([lv.solid for lv in levels], dtype=bool)
    L = len(lp)
    prox, qmin = float(P.contact_prox), float(P.q_min_prob or 0.0)
    entry_hist = np.zeros(H.N_ENTRY_CODES, dtype=np.int64)
    exit_hist = np.zeros(len(H.EXIT_CODES), dtype=np.int64)
    prev = np.zeros(n)
    last_dir = np.zeros(n, dtype=np.int8)             # 0 unknown, 1 up, 2 down
    cooldown_until = np.zeros(n, dtype=np.int64)
    inpos = np.zeros(n, dtype=bool)
    side = np.zeros(n, dtype=np.int8)
    stop = np.zeros(n); target = np.zeros(n); entry = np.zeros(n)
    touch = np.zeros((n, L), dtype=np.int16)
    latch = np.zeros((n, L), dtype=bool)
    path_pnl = np.zeros(n); sum_pnl2 = 0.0
    n_trades = n_wins = 0
    rows = np.arange(n)
    use_ml = lr is not None and not P.disable_ml_gate
    for i in range(steps):
        p = close[i].astype(np.float64)
        now = 60000 * i
        code = np.full(n, E["TOO_FAR"], dtype=np.int8)
        code[inpos] = E["IN_POSITION"]                                        # M:1508
        cool = ~inpos & (now < cooldown_until)
        code[cool] = E["COOLDOWN"]                                            # M:1516
        ev = ~inpos & ~cool
        ok = np.zeros(n, dtype=bool)
        if L == 0:
            code[ev] = E["NOLEVELS"]                                          # M:1524
        elif i == 0:
            code[ev] = E["DIR_UNKNOWN"]                                       # prev is None, M:1538
        else:
            d = np.where(p > prev + 1e-9, 1, np.where(p < prev - 1e-9, 2, last_dir)).astype(np.int8)   # M:1529-1537
            unk = ev & (d == 0)
            code[unk] = E["DIR_UNKNOWN"]
            idx = rows[ev & (d != 0)]
            if idx.size:
                dist_all = np.abs(lp[None, :] - p[idx, None])                  # M:1543: first minimum wins (argmin)
                k = np.argmin(dist_all, axis=1)
                dist = dist_all[np.arange(idx.size), k]
                near = dist <= prox
                idx, k, dist, dist_all = idx[near], k[near], dist[near], dist_all[near]      # the rest stay TOO_FAR
                if idx.size:
                    new = ~latch[idx, k]                                       # M:1562-1566
                    touch[idx, k] += new
                    keep = latch[idx] & (dist_all <= prox)                     # M:1572-1576
                    keep[np.arange(idx.size), k] = True
                    latch[idx] = keep
                    tc = touch[idx, k]
                    over = tc >= P.overtouch_limit                             # M:1579
                    base = np.maximum(0.0, 1.0 - (dist / max(0.0001, prox)))   # compute_confidence, M:1415-1427
                    base = base + np.where(solid[k], 0.08, 0.02)
                    base = base + np.where(tc <= 1, 0.10, np.where(tc == 2, -0.08, -(0.08 * 2)))
                    base = base + 0.03
                    conf = np.maximum(0.0, np.minimum(1.0, base))
                    low = ~over & (conf < qmin)                                # M:1628
                    up = d[idx] == 1
                    s_stop = np.where(up, lp[k] - P.stop_padding, lp[k] + P.stop_padding)       # M:1669-1674
                    s_tgt = np.where(up, lp[k] + P.tp_padding, lp[k] - P.tp_padding)
                    ml_low = np.zeros(idx.size, dtype=bool)
                    if use_ml:                                                 # M:1454-1466, z summed left to right
                        z = float(lr.intercept) + float(lr.coef[0]) * 1.0
                        z = z + float(lr.coef[1]) * np.abs(lp[k] - s_stop)
                        z = z + float(lr.coef[2]) * tc.astype(np.float64)
                        z = z + float(lr.coef[3]) * up.astype(np.float64)
                        ml_low = ~over & ~low & ~(1.0 / (1.0 + np.exp(-z)) >= qmin)
                    c2 = np.where(over, E["LEVEL_OVERTOUCHED"], np.where(low, E["CONF_LOW"], np.where(ml_low, E["ML_CONF_LOW"], E["OK"])))
                    code[idx] = c2
                    go = c2 == E["OK"]
                    gi = idx[go]
                    ok[gi] = True
                    side[gi] = np.where(up[go], 0, 1)
                    stop[gi] = s_stop[go]; target[gi] = s_tgt[go]
        if i > 0:                                                             # state update, M:2952-2955
            ch = p != prev
            last_dir[ch] = np.where(p[ch] > prev[ch], 1, 2)
        prev = p
        # exits for positions opened on an earlier tick (M:2966-3014), close-only
        xi = rows[inpos]
        if xi.size:
            px, lg = p[xi], side[xi] == 0
            is_stop = np.where(lg, px <= stop[xi], px >= stop[xi])
            is_tgt = ~is_stop & np.where(lg, px >= target[xi], px <= target[xi])
            out = is_stop | is_tgt
            oi = xi[out]
            pnl = np.where(lg[out], px[out] - entry[oi], entry[oi] - px[out])  # M:1926
            path_pnl[oi] += pnl
            sum_pnl2 += float(np.sum(pnl * pnl))
            n_trades += oi.size
            n_wins += int(np.sum(pnl > 0))
            exit_hist[X["STOP"]] += int(np.sum(is_stop))
            exit_hist[X["TARGET"]] += int(np.sum(is_tgt))
            cooldown_until[oi] = now + P.cooldown_s * 1000                     # M:269-270
            inpos[oi] = False
        gi = rows[ok]                                                          # opens (the exit test starts next tick)
        inpos[gi] = True
        entry[gi] = p[gi]
        entry_hist += np.bincount(code, minlength=H.N_ENTRY_CODES)
    n_open = int(inpos.sum())
    exit_hist[X["OPEN"]] += n_open
    return dict(entry_hist=entry_hist, exit_hist=exit_hist, n_trades=n_trades, n_wins=n_wins, n_open=n_open,
                sum_pnl=float(path_pnl.sum()), sum_pnl2=sum_pnl2, path_pnl=path_pnl)
