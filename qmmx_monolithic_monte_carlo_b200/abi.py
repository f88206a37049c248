"""ctypes mirrors of the POD structs and constants in include/qmmx_mc.h (the C ABI)."""
from __future__ import annotations

import ctypes as C

import numpy as np

ABI_VERSION = 1
MAX_LEVELS = 64
N_ENTRY_CODES = 20
N_EXIT_CODES = 4
PNL_HIST_BINS = 256
MAX_COMBOS = 1024
LR_DIM = 5
POLICY_HEADS = 3
POLICY_DIM = 7

OK, E_BADARG, E_CUDA, E_NOMEM, E_STATE, E_UNSUPPORTED, E_NCCL = 0, -1, -2, -3, -4, -5, -6

# reason codes of the reference (qmmx_monolithic.py:246-257) + policy_events sub-reasons
ENTRY_CODES = [
    "OK", "IN_POSITION", "COOLDOWN", "NOLEVELS", "DIR_UNKNOWN", "TOO_FAR",
    "LEVEL_OVERTOUCHED", "CONF_LOW", "ML_CONF_LOW", "COMBINED_LOW",
    "VETO_CONTRA_VOL_LONG", "VETO_CONTRA_VOL_SHORT", "VETO_EDGE_FATIGUE",
    "VETO_BOUNCE_BUDGET", "VETO_LEVEL_COOLDOWN", "ACC_BREAKOUT_GATE",
    "ONLINE_POLICY", "PRICE_STALE", "MISSING_API_KEY", "RISK_INVALID",
]
EXIT_CODES = ["STOP", "TARGET", "ESCALATE", "OPEN"]

F_ENABLE_VETO = 1 << 0
F_DISABLE_ML_GATE = 1 << 1
F_USE_BLEND = 1 << 2
F_DISABLE_POLICY_GATE = 1 << 3
F_EXIT_ON_HL = 1 << 4
F_TINY_PRICES = 1 << 5
F_ONLINE_RETRAIN = 1 << 6

GBM_ROUND_CENTS = 1 << 0

AGG_NU = 6 + N_ENTRY_CODES + N_EXIT_CODES + PNL_HIST_BINS
AGG_ND = 4
AGG_NX = 2
AGG_SLOTS = AGG_NU + AGG_ND + AGG_NX


class Params(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("flags", C.c_uint32),
        ("contact_prox", C.c_double), ("stop_padding", C.c_double), ("tp_padding", C.c_double),
        ("q_min_prob", C.c_double), ("w_rules", C.c_double), ("w_ml", C.c_double),
        ("veto_vol_strong", C.c_double), ("veto_prox", C.c_double),
        ("cooldown_s", C.c_int32), ("overtouch_limit", C.c_int32),
        ("steps_per_day", C.c_int32), ("reserved", C.c_int32),
    ]


class Gbm(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("flags", C.c_uint32),
        ("s0", C.c_float), ("mu_annual", C.c_float), ("sigma_annual", C.c_float), ("dt_years", C.c_float),
        ("sub_steps", C.c_int32), ("reserved", C.c_int32),
    ]


class Agg(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("combo", C.c_uint32),
        ("n_paths", C.c_uint64), ("n_steps", C.c_uint64), ("n_trades", C.c_uint64),
        ("n_wins", C.c_uint64), ("n_open", C.c_uint64), ("n_trade_overflow", C.c_uint64),
        ("entry_hist", C.c_uint64 * N_ENTRY_CODES), ("exit_hist", C.c_uint64 * N_EXIT_CODES),
        ("pnl_hist", C.c_uint64 * PNL_HIST_BINS),
        ("sum_pnl", C.c_double), ("sum_pnl2", C.c_double), ("sum_path_pnl2", C.c_double),
        ("sum_open_mtm", C.c_double), ("max_path_pnl", C.c_double), ("min_path_pnl", C.c_double),
        ("pnl_hist_lo", C.c_double), ("pnl_hist_hi", C.c_double),
    ]


TRADE_DTYPE = np.dtype([
    ("entry_bar", "<i4"), ("exit_bar", "<i4"), ("side", "u1"), ("reason", "u1"), ("level_id", "<u2"),
    ("entry", "<f4"), ("exit", "<f4"), ("pnl", "<f4"),
])
assert TRADE_DTYPE.itemsize == 24


def make_params(*, contact_prox=0.05, stop_padding=0.35, tp_padding=0.25, q_min_prob=0.60,
                cooldown_s=8, enable_veto=False, veto_vol_strong=0.25, veto_prox=0.06,
                disable_ml_gate=False, use_blend=False, w_rules=0.7, w_ml=0.3,
                disable_policy_gate=True, exit_on_hl=False, overtouch_limit=4,
                steps_per_day=390, tiny_prices=False, online_retrain=False) -> Params:
    """Engine settings with the reference's defaults (qmmx_monolithic.py:1371-1386, 1711-1715), with ONE deviation:
    `disable_policy_gate` defaults to True here, while the reference reads DISABLE_POLICY_GATE with default "0" (gate
    on, M:3090, M:2078).  The recorded run of the reference had `policy=False` and its shipped
    OnlinePolicy weights can never pass the gate (B.6: sigmoid = 0.37 < 0.60), so the reference default means "no trade
    ever"; pass disable_policy_gate=False together with set_policy(weights) to run the gate (float64 slow path)."""
    flags = ((F_ENABLE_VETO if enable_veto else 0) | (F_DISABLE_ML_GATE if disable_ml_gate else 0)
             | (F_USE_BLEND if use_blend else 0) | (F_DISABLE_POLICY_GATE if disable_policy_gate else 0)
             | (F_EXIT_ON_HL if exit_on_hl else 0) | (F_TINY_PRICES if tiny_prices else 0)
             | (F_ONLINE_RETRAIN if online_retrain else 0))
    return Params(C.sizeof(Params), flags, float(contact_prox), float(stop_padding), float(tp_padding),
                  float(q_min_prob), float(w_rules), float(w_ml), float(veto_vol_strong), float(veto_prox),
                  int(cooldown_s), int(overtouch_limit), int(steps_per_day), 0)


def make_gbm(*, s0=648.90, mu_annual=0.0, sigma_annual=0.15, dt_years=1.0 / (252.0 * 390.0),
             sub_steps=1, round_cents=False) -> Gbm:
    return Gbm(C.sizeof(Gbm), GBM_ROUND_CENTS if round_cents else 0, s0, mu_annual, sigma_annual,
               dt_years, sub_steps, 0)
