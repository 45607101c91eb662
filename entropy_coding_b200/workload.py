"""Seeded synthetic transform-unit streams.

The reference's own test inputs (three QCIF .y4m clips) are not available, so
every measured workload is synthetic: frames are tiled into CTUs, every CTU is
tiled with transform units, and coefficients are rounded Laplacians whose scale
decays with frequency.  Three scale presets mimic the densities the survey
probed on the reference encoder ("qp0" ~5 bit/sample, "qp32", "qp51").

Host-side numpy only; no device work happens here.
"""
from __future__ import annotations

import dataclasses

import numpy as np

from .types import (PARAMS_DT, SUB_ALIGN, SUB_DEP_QUANT, SUB_SIGN_HIDING, SUB_TERMINATE, TU_DT, TU_TS,
                    TU_TS_ALLOWED)

# (Laplace scale at DC, decay rate over normalised frequency, slice QP)
PRESETS = {
    "qp0": (48.0, 2.2, 0),
    "qp32": (3.0, 5.0, 32),
    "qp51": (1.1, 9.0, 51),
}

LUMA_SHAPES = [(4, 4), (8, 8), (16, 16), (32, 32), (8, 4), (4, 8), (16, 8), (8, 16), (32, 16), (16, 32),
               (16, 4), (4, 16), (64, 64), (32, 8), (8, 32)]


@dataclasses.dataclass
class Workload:
    tus: np.ndarray          # TU_DT[n_tu]
    tu_offsets: np.ndarray   # uint64[n_sub + 1]
    params: np.ndarray       # PARAMS_DT[n_sub]
    coeffs: np.ndarray       # int32 slab
    name: str = ""

    @property
    def n_sub(self) -> int:
        return len(self.params)

    @property
    def n_coeff_positions(self) -> int:
        return int(self.coeffs.size)

    def out_capacity(self, bytes_per_coeff: float = 2.5, slack: int = 256) -> np.ndarray:
        """Per-substream output capacity offsets (uint64[n_sub+1]), 16-byte aligned."""
        area = (1 << (self.tus["log2_w"].astype(np.int64) + self.tus["log2_h"].astype(np.int64)))
        csum = np.concatenate([[0], np.cumsum(area)])
        per_sub = csum[self.tu_offsets[1:].astype(np.int64)] - csum[self.tu_offsets[:-1].astype(np.int64)]
        cap = (np.ceil(per_sub * bytes_per_coeff).astype(np.int64) + slack + 15) & ~15
        return np.concatenate([[0], np.cumsum(cap)]).astype(np.uint64)

    def subset(self, idx) -> "Workload":
        """Substreams idx (in that order) re-based onto their own slab."""
        idx = np.asarray(idx, np.int64)
        tus_l, offs, co_l = [], [0], []
        base = 0
        for i in idx:
            a, b = int(self.tu_offsets[i]), int(self.tu_offsets[i + 1])
            t = self.tus[a:b].copy()
            if b > a:
                lo = int(t["coeff_offset"][0])
                last = t[-1]
                hi = int(last["coeff_offset"]) + (1 << (int(last["log2_w"]) + int(last["log2_h"])))
                co_l.append(self.coeffs[lo:hi])
                t["coeff_offset"] = t["coeff_offset"] - lo + base
                base += hi - lo
            tus_l.append(t)
            offs.append(offs[-1] + (b - a))
        return Workload(np.concatenate(tus_l) if tus_l else np.zeros(0, TU_DT), np.asarray(offs, np.uint64),
                        self.params[idx].copy(),
                        np.concatenate(co_l) if co_l else np.zeros(0, np.int32), self.name + "-subset")


def laplace_block(rng: np.random.Generator, n: int, w: int, h: int, scale: float, decay: float) -> np.ndarray:
    """n blocks (n, h, w) of rounded Laplacians with frequency-decaying scale."""
    y, x = np.mgrid[0:h, 0:w]
    f = (x / max(w, 1) + y / max(h, 1)) * 0.5
    s = (scale * np.exp(-decay * f)).astype(np.float32)
    if w > 32 or h > 32:  # zero-out region of 64-point transforms (JVET_C0024_ZERO_OUT_TH)
        s = np.where((x < 32) & (y < 32), s, 0.0).astype(np.float32)
    v = rng.laplace(0.0, 1.0, size=(n, h, w)).astype(np.float32) * s
    c = np.rint(v).astype(np.int32)
    np.clip(c, -32768, 32767, out=c)
    # residual_coding is only called for coded blocks (cbf = 1)
    empty = ~c.reshape(n, -1).any(axis=1)
    if empty.any():
        c[empty, 0, 0] = rng.choice(np.array([-2, -1, 1, 2], np.int32), size=int(empty.sum()))
    return c


def _emit(tus, coeffs_list, cursor, blocks, log2w, log2h, comp, flags):
    n = blocks.shape[0]
    area = 1 << (log2w + log2h)
    t = np.zeros(n, TU_DT)
    t["coeff_offset"] = cursor + np.arange(n, dtype=np.uint64) * area
    t["log2_w"], t["log2_h"], t["comp_id"], t["flags"] = log2w, log2h, comp, flags
    tus.append(t)
    coeffs_list.append(blocks.reshape(-1))
    return cursor + n * area


def random_substreams(n_sub: int, seed: int, preset: str | None = None, tus_per_sub=(4, 24),
                      ts_fraction: float = 0.1, flags: int | None = None, shapes=None,
                      chroma: bool = True, bdpcm_fraction: float = 0.0, sbt_fraction: float = 0.0) -> Workload:
    """Independent substreams of randomly shaped TUs (parity tests, decode sweep)."""
    from .types import TU_BDPCM, TU_SBT_ZEROOUT
    rng = np.random.default_rng(seed)
    shapes = shapes or LUMA_SHAPES
    names = list(PRESETS)
    tus, co, offs, cursor = [], [], [0], 0
    params = np.zeros(n_sub, PARAMS_DT)
    for i in range(n_sub):
        pn = preset or names[int(rng.integers(len(names)))]
        scale, decay, qp = PRESETS[pn]
        params["qp"][i] = qp
        params["init_id"][i] = 2 if flags is None else int(rng.integers(3))
        params["flags"][i] = (SUB_DEP_QUANT | SUB_TERMINATE | SUB_ALIGN) if flags is None else flags
        ntu = int(rng.integers(tus_per_sub[0], tus_per_sub[1] + 1))
        cnt = 0
        for _ in range(ntu):
            w, h = shapes[int(rng.integers(len(shapes)))]
            comp = int(rng.integers(3)) if chroma else 0
            if comp and (w > 32 or h > 32):
                w, h = min(w, 32), min(h, 32)
            f = 0
            if w <= 32 and h <= 32:
                f |= TU_TS_ALLOWED
                if rng.random() < ts_fraction:
                    f |= TU_TS
            if rng.random() < bdpcm_fraction and w <= 32 and h <= 32:
                f = TU_TS | TU_BDPCM  # bdpcm implies transform skip, flag not coded
            elif comp == 0 and rng.random() < sbt_fraction:
                f = TU_SBT_ZEROOUT    # SBT disables the ts flag
            blk = laplace_block(rng, 1, w, h, scale, 0.0 if (f & TU_TS) else decay)
            if f & TU_SBT_ZEROOUT:
                if w == 32:
                    blk[:, :, 16:] = 0
                if h == 32:
                    blk[:, 16:, :] = 0
                if not blk.any():
                    blk[0, 0, 0] = 1
            cursor = _emit(tus, co, cursor, blk, int(np.log2(w)), int(np.log2(h)), comp, f)
            cnt += 1
        offs.append(offs[-1] + cnt)
    return Workload(np.concatenate(tus), np.asarray(offs, np.uint64), params, np.concatenate(co),
                    f"random-{n_sub}x{tus_per_sub}-seed{seed}")


def frames_as_ctu_substreams(n_frames: int, width: int, height: int, seed0: int = 1000,
                             presets=("qp0", "qp32", "qp51"), ctu: int = 128, ts_fraction: float = 0.1,
                             per_frame: bool = False) -> Workload:
    """All-intra 4:2:0 frames; every CTU (or every frame when per_frame) is one substream.

    Luma TUs tile each CTU with one shape drawn per CTU; Cb/Cr follow at half size.
    The preset is drawn per substream (seed = seed0 + frame).
    """
    tus, co, offs, cursor = [], [], [0], 0
    plist = []
    names = list(presets)
    for fr in range(n_frames):
        rng = np.random.default_rng(seed0 + fr)
        frame_preset = names[int(rng.integers(len(names)))]
        cnt_frame = 0
        for cy in range(0, height, ctu):
            for cx in range(0, width, ctu):
                cw, ch = min(ctu, width - cx), min(ctu, height - cy)
                pn = frame_preset if per_frame else names[int(rng.integers(len(names)))]
                scale, decay, qp = PRESETS[pn]
                cand = [(w, h) for (w, h) in LUMA_SHAPES if cw % w == 0 and ch % h == 0 and w >= 4 and h >= 4]
                w, h = cand[int(rng.integers(len(cand)))]
                n = (cw // w) * (ch // h)
                ts = (rng.random() < ts_fraction) and w <= 32 and h <= 32
                f = (TU_TS_ALLOWED if (w <= 32 and h <= 32) else 0) | (TU_TS if ts else 0)
                cnt = 0
                blk = laplace_block(rng, n, w, h, scale, 0.0 if ts else decay)
                cursor = _emit(tus, co, cursor, blk, int(np.log2(w)), int(np.log2(h)), 0, f)
                cnt += n
                wc, hc = max(w // 2, 2), max(h // 2, 2)
                nc = (cw // 2 // wc) * (ch // 2 // hc)
                fc = TU_TS_ALLOWED if (wc <= 32 and hc <= 32) else 0
                for comp in (1, 2):
                    blk = laplace_block(rng, nc, wc, hc, scale * 0.6, decay)
                    cursor = _emit(tus, co, cursor, blk, int(np.log2(wc)), int(np.log2(hc)), comp, fc)
                    cnt += nc
                cnt_frame += cnt
                if not per_frame:
                    offs.append(offs[-1] + cnt)
                    plist.append((qp, 2, SUB_DEP_QUANT | SUB_TERMINATE | SUB_ALIGN))
        if per_frame:
            offs.append(offs[-1] + cnt_frame)
            plist.append((PRESETS[frame_preset][2], 2, SUB_DEP_QUANT | SUB_TERMINATE | SUB_ALIGN))
    params = np.zeros(len(plist), PARAMS_DT)
    for i, (qp, iid, fl) in enumerate(plist):
        params[i] = (qp, iid, fl, 0)
    return Workload(np.concatenate(tus), np.asarray(offs, np.uint64), params, np.concatenate(co),
                    f"{width}x{height}-AI-{n_frames}f-{'frame' if per_frame else 'ctu'}-substreams")


def frames_as_ctu_rows(n_frames: int, width: int, height: int, seed0: int = 1000, ctu: int = 128, **kw):
    """WPP layout (BASELINE.json configs[3]): one substream per CTU ROW, rows of a frame consecutive.

    Returns (workload, rows_per_frame, first_ctu_tus) where first_ctu_tus[i] is the number of TUs of
    row i's first CTU - the point after which the row below picks up its contexts
    (VTM EncSlice.cpp:2085-2089).  The slice QP of a frame is that of its first CTU.
    """
    wl = frames_as_ctu_substreams(n_frames, width, height, seed0=seed0, ctu=ctu, **kw)
    cpr = (width + ctu - 1) // ctu
    rows = (height + ctu - 1) // ctu
    assert wl.n_sub == n_frames * rows * cpr
    off = wl.tu_offsets.astype(np.int64)
    first = (off[1::cpr] - off[0::cpr][:len(off[1::cpr])]).astype(np.int64)
    params = wl.params[0::cpr].copy()
    for f in range(n_frames):  # one slice QP per frame
        params[f * rows:(f + 1) * rows] = params[f * rows]
    return (Workload(wl.tus, wl.tu_offsets[0::cpr].copy(), params, wl.coeffs,
                     f"{width}x{height}-WPP-{n_frames}f-row-substreams"), rows, first)


# ------------------------------------------------------------------------------------------------
# Batch shaping helpers for bench.py (host-side index arithmetic only)
# ------------------------------------------------------------------------------------------------
def merge_substreams(wl: Workload, group: int) -> Workload:
    """Every `group` consecutive substreams become one (same TUs, same slab): the 135 CTUs of a 1080p
    frame as ONE slice (BASELINE configs[2] read literally).  The merged substream keeps the
    parameters of its first member."""
    assert wl.n_sub % group == 0
    return Workload(wl.tus, wl.tu_offsets[::group].copy(), wl.params[::group].copy(), wl.coeffs,
                    wl.name + f"-merged{group}")


def tile_descriptors(wl: Workload, tiles: int, cap: np.ndarray):
    """Descriptor arrays of `tiles` back-to-back copies of a workload (copy k owns coefficient tile k,
    output slab k): (tus, tu_offsets, params, cap_offsets).  The coefficient slab itself is replicated on
    the device by the caller."""
    nco, ntu = int(wl.coeffs.size), len(wl.tus)
    tus = np.tile(wl.tus, tiles)
    tus["coeff_offset"] += np.repeat(np.arange(tiles, dtype=np.uint64) * np.uint64(nco), ntu)
    tu_off = np.concatenate([[0], (np.arange(tiles)[:, None] * ntu +
                                   wl.tu_offsets[1:].astype(np.int64)[None, :]).ravel()]).astype(np.uint64)
    cap_t = np.concatenate([[0], (np.arange(tiles)[:, None] * int(cap[-1]) +
                                  cap[1:].astype(np.int64)[None, :]).ravel()]).astype(np.uint64)
    return tus, tu_off, np.tile(wl.params, tiles), cap_t


def sweep_substreams(n_sub: int, seed: int, tus_per_sub=(64, 256)) -> Workload:
    """BASELINE configs[4] (decode sweep): independent substreams of 64+ small TUs each (count log-uniform
    in tus_per_sub, one block shape and one QP preset per substream, luma + chroma mixed), generated one
    vectorised draw per substream."""
    rng = np.random.default_rng(seed)
    shapes = [(4, 4), (8, 8), (8, 4), (4, 8), (16, 16), (16, 8), (8, 16), (4, 16), (16, 4)]
    weights = np.array([6, 5, 3, 3, 1, 1, 1, 1, 1], np.float64)
    weights /= weights.sum()
    names = list(PRESETS)
    tus, co, offs, cursor = [], [], [0], 0
    params = np.zeros(n_sub, PARAMS_DT)
    lo, hi = np.log(tus_per_sub[0]), np.log(tus_per_sub[1] + 1)
    for i in range(n_sub):
        scale, decay, qp = PRESETS[names[int(rng.integers(len(names)))]]
        params[i] = (qp, 2, SUB_DEP_QUANT | SUB_TERMINATE | SUB_ALIGN, 0)
        n = int(np.exp(rng.uniform(lo, hi)))
        w, h = shapes[int(rng.choice(len(shapes), p=weights))]
        comp = int(rng.integers(3))
        blk = laplace_block(rng, n, w, h, scale * (1.0 if comp == 0 else 0.6), decay)
        cursor = _emit(tus, co, cursor, blk, int(np.log2(w)), int(np.log2(h)), comp, TU_TS_ALLOWED)
        offs.append(offs[-1] + n)
    return Workload(np.concatenate(tus), np.asarray(offs, np.uint64), params, np.concatenate(co),
                    f"sweep-{n_sub}x{tus_per_sub}-seed{seed}")
