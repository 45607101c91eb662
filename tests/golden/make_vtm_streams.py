#!/usr/bin/env python
"""Generates tests/golden/vtm_*.npz: what the reference's slice writer REALLY emits.

TEST INFRASTRUCTURE; runs only in the build container (needs /root/reference, cmake, ninja).

Recipe (stand-in for BASELINE configs[0-1]; the three QCIF .y4m clips of test/hashes.txt are
absent from the checkout, `.MISSING_LARGE_BLOBS`):
  1. copy /root/reference to a scratch tree (VTM's CMake writes bin/ and lib/ into its source dir),
  2. apply tests/golden/vtm_record/vtm_record.patch: the final-pass bin encoder of
     VTM/Lib/EncoderLib/CABACWriter.h becomes vtm_record::RecordingBinEncoder (a subclass of the
     library's own BinEncoder_Std that forwards every call, so str.bin is unchanged - checked with
     --verify-unpatched) and EntropyCoding::CABACWriter::residual_coding / coding_tree_unit get a
     TU bracket / a CTU mark,
  3. build EncoderApp with STANDALONE_ENTROPY_CODEC=ON (CMakeLists.txt:8),
  4. encode a seeded synthetic QCIF clip (176x144 4:2:0 8 bit, 10 frames, TemporalSubsampleRatio 8
     -> 2 coded frames) with test/encoder_intra_vtm_qp{51,32,0}.cfg and the ctest command line
     (CMakeLists.txt:62), plus one WPP run (CTU 64, WaveFrontSynchro=1: 3 CTU rows per frame),
  5. turn the records into fixtures; each substream is first re-encoded with the compiled reference
     library (oracle/_ref) from the flat record - if that does not reproduce the payload the
     descriptor is missing something and the script stops.

    python tests/golden/make_vtm_streams.py [--work /tmp/ec_vtm] [--only qp51,qp32] [--skip-build]
"""
from __future__ import annotations

import argparse
import hashlib
import os
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from vtm_streams import parse_records, save_fixture  # noqa: E402

REF = "/root/reference"
CXX_FLAGS = ("-include cstdint -Wno-deprecated-declarations -Wno-array-compare -Wno-address -Wno-error "
             f"-I{ROOT}/include -I{ROOT}/entropy_coding_b200/host -I{HERE}/vtm_record")
CMDLINE = "-fr 60 -f 10 -wdt 176 -hgt 144 --InputChromaFormat=420 --InputBitDepth=8".split()
# A larger picture for the CTU-level parser (tests/test_syntax.py): 416x240, one coded frame, QP 27 - eight 128x128
# CTUs of which three are complete, so interior split decisions, more CU shapes and a realistic bin density.
CMDLINE_WQVGA = "-fr 60 -f 1 -wdt 416 -hgt 240 --InputChromaFormat=420 --InputBitDepth=8".split()
RUNS = {
    "wqvga_qp27": ("encoder_intra_vtm_qp32.cfg", ["-q", "27"]),
    # tools the three cfgs leave off but the CTU-level parser implements: cu_qp_delta (perceptual QP adaptation
    # turns pps.useDQP on, one quantisation group per CTU) and BDPCM (intra_bdpcm_*_flag, the cbf contexts under BDPCM)
    "dqp_bdpcm_qp32": ("encoder_intra_vtm_qp32.cfg", ["--PerceptQPA=1", "--BDPCM=1"]),  # (QPA stays slice-level at QCIF size: BDPCM + ALF chroma only)
    # cu_qp_delta for real: CU-level delta-QP search, quantisation groups two subdivisions below the CTU
    "cudqp_qp32": ("encoder_intra_vtm_qp32.cfg", ["--MaxDeltaQP=1", "--MaxCuDQPSubdiv=2"]),
    # cu_chroma_qp_offset: with offset lists in the PPS the VTM encoder cycles cu_chroma_qp_offset_idx over the chroma
    # quantisation groups (EncCu::xCompressCU); a lower QP so that the loop-filter search may also pick CC-ALF
    "cqpadj_qp27": ("encoder_intra_vtm_qp32.cfg", ["-q", "27", "--MaxCuChromaQpOffsetSubdiv=2", "--CbQpOffsetList=2,-2",
                                                   "--CrQpOffsetList=-2,2", "--CbCrQpOffsetList=1,-1"]),
    # CC-ALF: a clip whose chroma texture follows the luma texture (synth_qcif(chroma_detail=True)), two frames
    "ccalf_qp37": ("encoder_intra_vtm_qp32.cfg", ["-q", "37", "-f", "2"]),
    "ccalf_qp27": ("encoder_intra_vtm_qp32.cfg", ["-q", "27", "-f", "2"]),
    # chroma BDPCM: a clip with a checker patch in the chroma planes (synth_qcif(chroma_checker=True))
    "cbdpcm_qp27": ("encoder_intra_vtm_qp32.cfg", ["-q", "27", "-f", "2", "--BDPCM=1"]),
    "qp51": ("encoder_intra_vtm_qp51.cfg", []),
    "qp32": ("encoder_intra_vtm_qp32.cfg", []),
    "qp0": ("encoder_intra_vtm_qp0.cfg", []),
    # WPP: one substream per CTU row, contexts handed over after the first CTU of the row above
    "wpp_qp32": ("encoder_intra_vtm_qp32.cfg",
                 ["--CTUSize=64", "--MaxBTNonISlice=64", "--MaxTTNonISlice=64", "--WaveFrontSynchro=1"]),
}


def synth_qcif(seed=2024, w=176, h=144, frames=10, chroma_detail=False, chroma_checker=False) -> bytes:
    """Seeded clip with gradients, a moving blob, a checker patch (sharp edges: transform skip) and
    noise, so that the intra search uses a mix of tools."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    out = bytearray()
    tex = rng.normal(0, 1, (h, w))
    for f in range(frames):
        y = 90 + 0.5 * xx + 0.3 * yy
        y += 40 * np.sin((xx + 3 * f) / 9.0) * np.cos(yy / 13.0)
        cx, cy = 60 + 4 * f, 70 + 2 * f
        y += 60 * np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / 300.0)
        y[100:130, 20:90] = 30 + 180 * (((xx[100:130, 20:90].astype(int) // 3) +
                                          (yy[100:130, 20:90].astype(int) // 5)) % 2)
        y += 6 * tex + rng.normal(0, 2, (h, w))
        if chroma_detail:  # fine luma structure the chroma planes share (below)
            y += 25 * np.sin((xx + 3 * f) / 2.5) * np.cos(yy / 3.0)
        u = 128 + 30 * np.sin(xx[::2, ::2] / 20.0 + f / 5.0) + rng.normal(0, 1.5, (h // 2, w // 2))
        v = 128 + 30 * np.cos(yy[::2, ::2] / 17.0) + 20 * (xx[::2, ::2] > 88) + rng.normal(0, 1.5, (h // 2, w // 2))
        if chroma_detail:  # chroma texture that follows the luma texture: what the cross-component ALF predicts
            d = (6 * tex + 40 * np.sin((xx + 3 * f) / 2.5) * np.cos(yy / 3.0))[::2, ::2]
            u = u + 0.8 * d
            v = v - 0.6 * d
        if chroma_checker:  # sharp horizontal / vertical chroma edges: what chroma BDPCM / transform skip are for
            cxx, cyy = xx[::2, ::2].astype(int), yy[::2, ::2].astype(int)
            u[8:40, 8:56] = 60 + 140 * (((cxx[8:40, 8:56] // 4) + (cyy[8:40, 8:56] // 3)) % 2)
            v[30:64, 40:80] = 200 - 150 * ((cxx[30:64, 40:80] // 5) % 2)
        for p in (y, u, v):
            out += np.clip(np.rint(p), 0, 255).astype(np.uint8).tobytes()
    return bytes(out)


def sh(cmd, **kw):
    print("[vtm]", " ".join(cmd) if isinstance(cmd, list) else cmd, flush=True)
    subprocess.check_call(cmd, **kw)


def build(work: str, patched: bool) -> str:
    tree = os.path.join(work, "tree_patched" if patched else "tree_orig")
    exe = os.path.join(tree, "bin", "EncoderAppStatic")
    if os.path.exists(exe):
        return exe
    if not os.path.exists(tree):
        shutil.copytree(REF, tree, symlinks=True)
        if patched:
            sh(["patch", "-p1", "-i", os.path.join(HERE, "vtm_record", "vtm_record.patch")], cwd=tree)
    bdir = os.path.join(tree, "build")
    os.makedirs(bdir, exist_ok=True)
    sh(["cmake", "..", "-G", "Ninja", "-DCMAKE_BUILD_TYPE=Release", "-DSTANDALONE_ENTROPY_CODEC=ON",
        "-DENABLE_LOGGING=OFF", f"-DCMAKE_CXX_FLAGS={CXX_FLAGS}"], cwd=bdir, stdout=subprocess.DEVNULL)
    sh(["ninja", "-j", str(os.cpu_count() or 4), "EncoderApp"], cwd=bdir, stdout=subprocess.DEVNULL)
    return exe


def cmdline_for(name: str) -> list[str]:
    return CMDLINE_WQVGA if name.startswith("wqvga") else CMDLINE


def encode(exe: str, rundir: str, clip: str, cfg: str, extra: list[str], record: bool, cmdline=None) -> bytes:
    os.makedirs(rundir, exist_ok=True)
    rec = os.path.join(rundir, "rec.bin")
    if record and os.path.exists(rec):
        os.remove(rec)
    env = dict(os.environ)
    if record:
        env["VTM_RECORD_OUT"] = rec
    with open(os.path.join(rundir, "enc.log"), "w") as log:
        sh([exe, "-i", clip, "-c", os.path.join(REF, "test", cfg)] + (cmdline or CMDLINE) + extra, cwd=rundir, env=env,
           stdout=log, stderr=subprocess.STDOUT)
    return open(os.path.join(rundir, "str.bin"), "rb").read()


def to_fixture(name: str, rundir: str, cfg: str, extra: list[str], cmdline=None) -> None:
    from oracle.checker import SUB_ALIGN, SUB_TERMINATE, Checker
    ref = Checker("ref")
    subs = parse_records(os.path.join(rundir, "rec.bin"))
    strbin = np.fromfile(os.path.join(rundir, "str.bin"), np.uint8)
    est = []
    for i, s in enumerate(subs):
        assert int(s["cmds"][-1]) == ((3 << 29) | 1), "substream must end with encodeBinTrm(1) (end_of_slice)"
        init, _ = ref.ctx_init(s["qp"], s["init_id"])
        fresh = (init["state0"] == s["ctx_in"][:, 0]).all() and (init["state1"] == s["ctx_in"][:, 1]).all()
        ctx_in = None
        if not fresh:
            ctx_in = np.zeros(379, init.dtype)
            ctx_in["state0"], ctx_in["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
        out, nb, _ = ref.encode_stream(s["cmds"][:-1], s["tus"], s["coeffs"], s["qp"], s["init_id"],
                                       s["flags"] | SUB_TERMINATE | SUB_ALIGN, ctx_in=ctx_in)
        assert out.tobytes() == s["payload"].tobytes() and nb == s["num_bins"], \
            f"{name} substream {i}: the flat record does not reproduce the writer's bytes"
        # BitEstimator_Std over the substream's TU list from fresh contexts (the RDO view of the same blocks)
        if len(s["tus"]):
            tot, per_tu = ref.estimate_residual(s["tus"], s["coeffs"], s["qp"], s["init_id"], s["flags"],
                                                use_mts=bool(s["tools"] & 4))
            est.append(np.concatenate([[tot], per_tu]).astype(np.uint64))
        else:
            est.append(np.zeros(1, np.uint64))
    extra_arrays = {"str_bin": strbin,
                    "str_bin_md5": np.frombuffer(hashlib.md5(strbin.tobytes()).hexdigest().encode(), np.uint8),
                    "cmdline": np.frombuffer((cfg + " " + " ".join((cmdline or CMDLINE) + extra)).encode(), np.uint8)}
    for i, e in enumerate(est):
        extra_arrays[f"est{i}"] = e
    path = os.path.join(HERE, f"vtm_{name}.npz")
    save_fixture(path, subs, extra_arrays)
    tot_bins = sum(s["num_bins"] for s in subs)
    print(f"[vtm] {path}: {len(subs)} substreams, {sum(len(s['tus']) for s in subs)} TUs, {tot_bins} bins, "
          f"{sum(s['payload'].size for s in subs)} payload bytes, str.bin md5 "
          f"{hashlib.md5(strbin.tobytes()).hexdigest()}, {os.path.getsize(path)} bytes", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--work", default="/tmp/ec_vtm")
    ap.add_argument("--only", default=",".join(RUNS))  # qp0 and wqvga_qp27 take tens of minutes each
    ap.add_argument("--skip-encode", action="store_true", help="reuse <work>/<run>/rec.bin + str.bin")
    ap.add_argument("--verify-unpatched", action="store_true",
                    help="also build the unpatched tree and check str.bin is identical (qp51 only: minutes)")
    ap.add_argument("--exe", default=None, help="use this patched EncoderApp instead of building one")
    a = ap.parse_args()
    os.makedirs(a.work, exist_ok=True)
    clip = os.path.join(a.work, "synth_qcif.yuv")
    open(clip, "wb").write(synth_qcif())
    clip_wqvga = os.path.join(a.work, "synth_wqvga.yuv")
    open(clip_wqvga, "wb").write(synth_qcif(seed=2025, w=416, h=240, frames=1))
    clip_cd = os.path.join(a.work, "synth_qcif_cd.yuv")
    open(clip_cd, "wb").write(synth_qcif(seed=2026, chroma_detail=True))
    clip_cc = os.path.join(a.work, "synth_qcif_cc.yuv")
    open(clip_cc, "wb").write(synth_qcif(seed=2027, chroma_checker=True))
    exe = a.exe or (None if a.skip_encode else build(a.work, True))
    for name in a.only.split(","):
        cfg, extra = RUNS[name]
        rundir = os.path.join(a.work, name)
        cl = cmdline_for(name)
        if not a.skip_encode:
            encode(exe, rundir, clip_wqvga if name.startswith("wqvga") else clip_cd if name.startswith("ccalf") else clip_cc if name.startswith("cbdpcm") else clip,
                   cfg, extra, True, cl)
        to_fixture(name, rundir, cfg, extra, cl)
    if a.verify_unpatched:
        orig = build(a.work, False)
        b0 = encode(orig, os.path.join(a.work, "qp51_orig"), clip, RUNS["qp51"][0], [], False)
        b1 = open(os.path.join(a.work, "qp51", "str.bin"), "rb").read()
        assert b0 == b1, "the recording patch changed str.bin"
        print("[vtm] unpatched EncoderApp writes the same str.bin:", hashlib.md5(b0).hexdigest())


if __name__ == "__main__":
    main()
