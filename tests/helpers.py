"""Shared helpers for the parity tests (host-side only)."""
import numpy as np

from entropy_coding_b200.types import (CTX_DT, PARAMS_DT, SUB_ALIGN, SUB_TERMINATE, mk_align, mk_ctx, mk_ep, mk_rem,
                                       mk_trm)

T = SUB_TERMINATE | SUB_ALIGN


def random_cmds(rng, n, with_align=False):
    cmds = []
    p1 = rng.random()
    for _ in range(n):
        k = rng.random()
        if k < 0.6:
            cmds.append(mk_ctx(int(rng.integers(0, 379)), int(rng.random() < p1)))
        elif k < 0.8:
            nb = int(rng.integers(0, 25))
            cmds.append(mk_ep(int(rng.integers(0, 1 << nb)) if nb else 0, nb))
        elif k < 0.93:
            rice = int(rng.integers(0, 5))
            vmax = ((4100 + (1 << (15 - rice))) << rice) - 1
            cmds.append(mk_rem(min(int(rng.integers(0, 1 << int(rng.integers(1, 18)))), vmax), rice))
        elif k < 0.97 or not with_align:
            cmds.append(mk_trm(0))
        else:
            cmds.append(mk_align())
    return np.asarray(cmds, np.uint32)


def mixed_stream(rng, n_tu, raw_per_gap=(0, 40)):
    """Recorded commands (non-residual ctx ids 0..85, EP, REM, trm(0)) with one TU marker per TU."""
    from entropy_coding_b200.types import mk_tu
    cmds = []
    for _ in range(n_tu):
        for _ in range(int(rng.integers(raw_per_gap[0], raw_per_gap[1] + 1))):
            k = rng.random()
            if k < 0.7:
                cmds.append(mk_ctx(int(rng.integers(0, 86)), int(rng.random() < 0.3)))
            elif k < 0.85:
                nb = int(rng.integers(1, 25))
                cmds.append(mk_ep(int(rng.integers(0, 1 << nb)), nb))
            elif k < 0.95:
                cmds.append(mk_rem(int(rng.integers(0, 3000)), int(rng.integers(0, 4))))
            else:
                cmds.append(mk_trm(0))
        cmds.append(mk_tu())
    for _ in range(int(rng.integers(0, 6))):
        cmds.append(mk_ctx(int(rng.integers(292, 357)), int(rng.integers(2))))
    return np.asarray(cmds, np.uint32)


def expected_values(cmds):
    """What decoding a stream must return for every command (no ALIGN inside)."""
    cmds = np.asarray(cmds, np.uint32)
    kind = cmds >> 29
    return np.where(kind == 0, (cmds >> 9) & 1, np.where((kind == 1) | (kind == 2), cmds & 0xFFFFFF, 0)).astype(
        np.uint32)


def ctx_pairs(ctx):
    """CTX_DT array -> (n, 2) uint16."""
    return np.stack([ctx["state0"], ctx["state1"]], axis=-1)


def params_of(qp, init_id, flags):
    p = np.zeros(1, PARAMS_DT)
    p["qp"], p["init_id"], p["flags"] = qp, init_id, flags
    return p


def sub_range(wl, i):
    a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
    lo = int(wl.tus[a]["coeff_offset"])
    last = wl.tus[b - 1]
    hi = int(last["coeff_offset"]) + (1 << (int(last["log2_w"]) + int(last["log2_h"])))
    return a, b, lo, hi


def oracle_encode_workload(chk, wl, want_ctx=False):
    """Runs a CPU checker over every substream of a workload."""
    res = []
    for i in range(wl.n_sub):
        a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
        p = wl.params[i]
        res.append(chk.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]),
                                       want_ctx=want_ctx))
    return res
