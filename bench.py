#!/usr/bin/env python
"""bench.py - CABAC encode/decode throughput of the B200 engine (BASELINE.json metric).

A "step" is one pass of the hot path over one batch: encode every substream of the
batch (residual binarisation + arithmetic coder + substream packing) and decode the
produced bitstreams back to coefficients.  Workload at every N: BASELINE.json
configs[2], "synthetic 1080p all-intra batch, 64 frames", one substream per 128x128
CTU (135 per frame, 8 640 per GPU); each rank codes its own 64 frames (weak scaling,
no data-path collective - SURVEY.md section 8e).

  value   Mbins/s with inputs resident in HBM (CUDA events on the launching stream,
          max over ranks), bins = BinCounter::getAll() of encode + decode
  e2e     same metric through the host-pointer C ABI (pinned host buffers, H2D of the
          coefficient slab / bitstreams and D2H of the results inside the timed region)
  roofline / issue_roofline   dominant kernel vs measured HBM copy bandwidth, and vs the
          per-SM integer issue bound (north_star)
  cpu_baseline  the compiled reference (oracle/_ref) - or the C port when it is absent -
          on this box's host cores over a bounded sample of the same workload

`--impl reference` times only the CPU reference with all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cabac_encode_decode_throughput"
UNIT = "Mbins/s"
WORKLOAD = "1080p-all-intra-64-frames, one substream per 128x128 CTU (8640 substreams/GPU), mixed qp0/qp32/qp51 presets"


_SAVED_STDOUT = None  # file descriptor 1 as the driver gave it, while a multi-rank run has it pointed at stderr


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=64, help="frames per GPU (64 = the BASELINE config)")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--substreams", default="ctu", choices=["ctu", "frame"],
                    help="ctu: one substream per 128x128 CTU (the bench line); frame: one per frame (64 long slices, latency-bound)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-sample-frames", type=int, default=16)
    ap.add_argument("--no-legs", action="store_true",
                    help="skip the short legs for the other BASELINE configs (1080p as 64 slices, 4K WPP, decode sweep)")
    ap.add_argument("--wpp-frames", type=int, default=256, help="4K WPP leg: frames of the whole job (sharded over the ranks)")
    ap.add_argument("--sweep-max", type=int, default=1_000_000, help="decode sweep leg: largest substream count per launch")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clock / throttle-reason samples during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # samples under load = the upper half (the sampler also sees the idle edges)
        hi = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": float(np.median(hi)), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- CPU reference leg
def cpu_checker():
    """The compiled reference when it travelled with the repo (oracle/_ref), else the C port."""
    from oracle import checker
    if checker.available("ref"):
        return checker.Checker("ref"), "reference"
    if not checker.available("oracle"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    return checker.Checker("oracle"), "port"


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def cpu_pass(chk, wl, threads):
    """One encode + decode pass of the CPU reference over a workload.  Returns (seconds, bins)."""
    cap = wl.out_capacity()
    out = np.zeros(int(cap[-1]), np.uint8)
    sizes = np.zeros(wl.n_sub, np.uint32)
    t_enc, bins, _ = chk.time_encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, threads, out, cap, sizes)
    dec = np.zeros(wl.coeffs.size, np.int32)
    t_dec = chk.time_decode_residual(out, cap, sizes, wl.tus, wl.tu_offsets, wl.params, dec, threads)
    assert (dec == wl.coeffs).all(), "CPU reference decode(encode(x)) != x"
    return t_enc, t_dec, bins


def run_reference_arm(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    from entropy_coding_b200 import workload as W
    chk, kind = cpu_checker()
    threads = host_threads()
    frames = max(1, min(args.cpu_sample_frames, args.frames))
    wl = W.frames_as_ctu_substreams(frames, args.width, args.height, seed0=1000, per_frame=args.substreams == "frame")
    for _ in range(max(1, args.warmup)):
        cpu_pass(chk, wl, threads)
    t_enc = t_dec = 0.0
    bins = 0
    for _ in range(args.steps):
        a, b, n = cpu_pass(chk, wl, threads)
        t_enc += a; t_dec += b; bins += 2 * n
    total = t_enc + t_dec
    val = bins / total / 1e6
    sample = (f"{frames} of {args.frames} frames ({wl.n_sub} substreams, {bins // (2 * args.steps)} bins) per step, "
              f"encode+decode, {threads} threads, dynamic substream hand-out")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "encode_mbins_s": bins / 2 / t_enc / 1e6, "decode_mbins_s": bins / 2 / t_dec / 1e6,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- B200 arm
def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_profile_facts():
    """ncu-derived per-launch facts committed under profiles/ (traffic, instructions per bin)."""
    p = os.path.join(ROOT, "profiles", "kernel_facts.json")
    if os.path.exists(p):
        try:
            return json.load(open(p))
        except Exception:
            pass
    return {}


def dev_bytes(t, a: np.ndarray):
    """numpy array (any dtype, incl. structured) -> device uint8 tensor."""
    import torch
    a = np.ascontiguousarray(a)
    return torch.from_numpy(a.view(np.uint8).reshape(-1)).to(t)


def bind_to_gpu_numa_node(local_rank: int) -> str:
    """One process per GPU: run (and first-touch the pinned host slabs) on the NUMA node the GPU hangs off,
    so the end-to-end leg's PCIe traffic does not cross the socket interconnect.  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        index = int(visible.split(",")[local_rank]) if visible and visible.replace(",", "").isdigit() else local_rank
        bdf = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
        if len(bdf.split(":")[0]) == 8:
            bdf = bdf[4:]
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return "single node"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return f"node {node}: none of its cores allowed"
        os.sched_setaffinity(0, allowed)
        return f"node {node}, {len(allowed)} cores"
    except Exception as e:  # noqa: BLE001 - placement is an optimisation, never a failure
        return f"unbound ({type(e).__name__})"


def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    import entropy_coding_b200 as E
    from entropy_coding_b200 import workload as W

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - entropy_coding_b200 has no CPU fallback")
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else "single process"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # stdout carries ONE JSON line: anything a library writes to file descriptor 1 while the job runs
        # (the "NCCL version ..." banner at communicator creation) goes to stderr; the descriptor is
        # restored for the result line
        global _SAVED_STDOUT
        sys.stdout.flush()
        _SAVED_STDOUT = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- workload: this rank's 64 frames (weak scaling: different frames per rank) ----
    wl = W.frames_as_ctu_substreams(args.frames, args.width, args.height, seed0=1000 + rank * args.frames,
                                    per_frame=args.substreams == "frame")
    n_sub, n_tu, n_co = wl.n_sub, len(wl.tus), int(wl.coeffs.size)
    cap = wl.out_capacity()
    eng = E.CabacEngine(local_rank)
    # a real (non-default) stream: kernels, copies and the timing events all go onto it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    eng.set_stream(stream.cuda_stream)

    d_tus, d_tuoff = dev_bytes(dev, wl.tus), dev_bytes(dev, wl.tu_offsets)
    d_par, d_cap = dev_bytes(dev, wl.params), dev_bytes(dev, cap)
    d_co = torch.from_numpy(wl.coeffs).to(dev)
    d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
    d_packed = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
    d_entry = torch.zeros(n_sub + 1, dtype=torch.int64, device=dev)
    d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_ebins = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_est = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_eflags = torch.zeros(n_tu, dtype=torch.uint8, device=dev)
    d_dec = torch.empty(n_co, dtype=torch.int32, device=dev)
    d_dbins = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_dbits = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_dst = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_dflags = torch.zeros(n_tu, dtype=torch.uint8, device=dev)

    ev = lambda: torch.cuda.Event(enable_timing=True)
    t_kernel = {"encode_residual_kernel": 0.0, "pack": 0.0, "decode_residual_kernel": 0.0}

    def step(timed):
        e = [ev() for _ in range(4)] if timed else None
        if timed: e[0].record(stream)
        eng.encode_residual_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_out, d_cap, d_sizes, d_ebins, d_eflags,
                                status=d_est)
        if timed: e[1].record(stream)
        eng.pack_dev(d_out, d_cap, d_sizes, n_sub, d_packed, d_entry)
        if timed: e[2].record(stream)
        eng.decode_residual_dev(d_packed, d_entry, d_tus, d_tuoff, d_par, n_sub, d_dec, d_dbins, d_dbits, d_dflags,
                                status=d_dst)
        if timed: e[3].record(stream)
        return e

    for _ in range(max(args.warmup, 3)):  # never fewer than 3 warm-up steps
        step(False)
    barrier()
    launches0 = eng.launch_count
    clocks = ClockSampler(local_rank)
    clocks.start()
    t0, t1 = ev(), ev()
    t0.record(stream)
    evs = [step(True) for _ in range(args.steps)]
    t1.record(stream)
    barrier()
    clk = clocks.stop()
    launches = eng.launch_count - launches0
    elapsed_ms = t0.elapsed_time(t1)
    for e in evs:
        t_kernel["encode_residual_kernel"] += e[0].elapsed_time(e[1])
        t_kernel["pack"] += e[1].elapsed_time(e[2])
        t_kernel["decode_residual_kernel"] += e[2].elapsed_time(e[3])
    for k in t_kernel:
        t_kernel[k] /= args.steps

    # ---- correctness of what was timed ----
    assert int(d_est.abs().max()) == 0 and int(d_dst.abs().max()) == 0, "substream status set"
    assert torch.equal(d_dec, d_co), "decode(encode(x)) != x on the timed batch"
    assert torch.equal(d_ebins, d_dbins), "encoder and decoder disagree on bin counts"
    enc_bins = int(d_ebins.to(torch.int64).sum())
    out_bytes = int(d_sizes.to(torch.int64).sum())
    bins_step = 2 * enc_bins

    tt = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    tb = torch.tensor([float(bins_step)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(tb, op=dist.ReduceOp.SUM)
    ms_per_step = float(tt) / args.steps
    value = float(tb) / (ms_per_step * 1e-3) / 1e6
    # ---- byte parity of the timed batch: every 64th substream against the CPU reference / oracle ----
    verified = "decode(encode(x)) == x and equal bin counts on the timed batch"
    if rank == 0:
        chk, kind = cpu_checker()
        h_entry = d_entry.cpu().numpy().astype(np.int64)
        h_packed = d_packed[:int(h_entry[-1])].cpu().numpy()
        step_k = 64 if args.substreams == "ctu" else 16
        picked = list(range(0, n_sub, step_k))
        for i in picked:
            a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
            pp = wl.params[i]
            eb = chk.encode_residual(wl.tus[a:b], wl.coeffs, int(pp["qp"]), int(pp["init_id"]), int(pp["flags"]))[0]
            assert h_packed[h_entry[i]:h_entry[i + 1]].tobytes() == eb.tobytes(), f"substream {i}: GPU bytes != CPU {kind}"
        verified += f"; packed bytes of every {step_k}th substream ({len(picked)}) == the CPU {kind} coder's"
    res = dict(wl=wl, cap=cap, eng=eng, value=value, ms_per_step=ms_per_step, t_kernel=t_kernel, clk=clk,
               launches=launches, enc_bins=enc_bins, out_bytes=out_bytes, n_sub=n_sub, n_tu=n_tu, n_co=n_co,
               world=world, rank=rank, dev=dev, barrier=barrier, all_bins=float(tb), numa=numa, verified=verified,
               stream=stream)
    if not args.no_legs and args.substreams == "ctu":
        # BASELINE configs[2] read literally: the same 64 frames, ONE substream (slice) per frame
        res["leg_slices"] = leg_slices(args, res, d_tus, d_co, d_out, d_packed, d_dec)
    # free the device-resident buffers before the end-to-end leg allocates its own
    del d_out, d_packed, d_dec, d_co
    torch.cuda.empty_cache()
    return res



# --------------------------------------------------------------------------- legs for the other BASELINE configs
def _sm_mhz(r):
    return r["clk"].get("sm_mhz") or 1965.0


def _fractions(r, kernel, bins, ms, alg_bytes):
    """HBM and issue fractions of one timed kernel pass (instructions per bin: the ncu capture of the bench
    batch, profiles/kernel_facts.json - the same kernel, so the per-bin cost carries over approximately)."""
    peak, _ = load_peaks()
    out = {"hbm_frac": alg_bytes / (ms * 1e-3) / 1e9 / peak}
    ipb = load_profile_facts().get(kernel, {}).get("warp_instr_per_bin")
    if ipb:
        out["issue_frac"] = (bins / (ms * 1e-3)) / (148 * 4 * _sm_mhz(r) * 1e6 / ipb)
        out["issue_kernel"] = kernel
    return out


def _time_passes(stream, fn, warm=1, reps=2):
    import torch
    for _ in range(warm):
        fn()
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1)
        best = t if best is None else min(best, t)
    return best


def leg_slices(args, r, d_tus, d_co, d_out, d_packed, d_dec):
    """configs[2] as literally written: 64 frames as 64 independent slices (one substream per frame).
    Same TUs and coefficients as the headline batch, merged 135 CTUs -> 1 slice; few, long substreams."""
    import torch
    from entropy_coding_b200 import workload as W
    eng, dev, stream = r["eng"], r["dev"], r["stream"]
    per_frame = r["n_sub"] // args.frames
    wl = W.merge_substreams(r["wl"], per_frame)
    n_sub = wl.n_sub
    cap = wl.out_capacity()
    d_tuoff, d_par, d_cap = dev_bytes(dev, wl.tu_offsets), dev_bytes(dev, wl.params), dev_bytes(dev, cap)
    d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_bins, d_st, d_dbins = torch.zeros_like(d_sizes), torch.zeros_like(d_sizes), torch.zeros_like(d_sizes)
    d_entry = torch.zeros(n_sub + 1, dtype=torch.int64, device=dev)

    def enc():
        eng.encode_residual_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_out, d_cap, d_sizes, d_bins, status=d_st)
        eng.pack_dev(d_out, d_cap, d_sizes, n_sub, d_packed, d_entry)

    def dec():
        eng.decode_residual_dev(d_packed, d_entry, d_tus, d_tuoff, d_par, n_sub, d_dec, d_dbins, status=d_st)

    l0 = eng.launch_count
    t_enc = _time_passes(stream, enc, warm=1, reps=1)
    t_dec = _time_passes(stream, dec, warm=0, reps=1)
    assert int(d_st.abs().max()) == 0 and torch.equal(d_dec, d_co) and torch.equal(d_bins, d_dbins), "slices leg round trip"
    bins = int(d_bins.to(torch.int64).sum())
    nbytes = int(d_sizes.to(torch.int64).sum())
    tb = torch.tensor([float(bins), t_enc, t_dec], dtype=torch.float64, device=dev)
    if r["world"] > 1:
        import torch.distributed as dist
        tmax = tb.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tb, op=dist.ReduceOp.SUM)
        t_enc, t_dec, bins_all = float(tmax[1]), float(tmax[2]), float(tb[0])
    else:
        bins_all = float(bins)
    alg = 4 * r["n_co"] + 16 * r["n_tu"] + nbytes
    leg = {"config": "BASELINE configs[2] as written: 1080p all-intra, 64 frames = 64 independent slices per GPU "
                     "(the headline batch's TUs, 135 CTUs merged per slice)",
           "substreams_per_gpu": n_sub, "bins_per_pass_per_gpu": bins, "bitstream_bytes_per_gpu": nbytes,
           "encode_mbins_s": bins_all / t_enc / 1e3, "decode_mbins_s": bins_all / t_dec / 1e3,
           "encode_ms": t_enc, "decode_ms": t_dec, "output_mb_s": nbytes * r["world"] / t_enc / 1e3,
           "encode": _fractions(r, "encode_residual_kernel", bins, t_enc, alg),
           "decode": _fractions(r, "decode_residual_kernel", bins, t_dec, alg),
           "gpu_launches": int(eng.launch_count - l0),
           "note": "few, long substreams: one serial chain per slice; decode has no intra-substream parallelism "
                   "(every bin selects the next context), so its figure is the latency floor of 64 chains",
           "decode_cycles_per_bin": t_dec * 1e-3 * _sm_mhz(r) * 1e6 / (bins / max(1, n_sub)),
           "decode_chain": "one warp per slice walks its bins one after another: decode_cycles_per_bin SM clocks per bin "
                           "(41 dependent warp instructions at ~5 clocks); a thread-per-slice kernel is slower still "
                           "(64 scalar instructions per bin, 418 vs 576 Mbins/s, profiles/r2_decoder_modes.md); a 3+ GHz "
                           "out-of-order CPU core does a bin in ~70 clocks, so 64 chains cannot beat 16 such cores - "
                           "throughput comes from more slices in flight (the headline batch, decode_sweep)",
           "verified": "decode(encode(x)) == x, equal bin counts"}
    if r["rank"] == 0 and not args.no_cpu_baseline:
        chk, kind = cpu_checker()
        threads = host_threads()
        k = max(1, min(threads, n_sub, args.cpu_sample_frames))
        sample = wl.subset(range(k))
        cpu_pass(chk, sample, threads)
        te, td, nb = cpu_pass(chk, sample, threads)
        leg["cpu_baseline"] = {"value": 2 * nb / (te + td) / 1e6, "unit": UNIT, "cores": min(threads, k), "kind": kind,
                               "encode_mbins_s": nb / te / 1e6, "decode_mbins_s": nb / td / 1e6,
                               "sample": f"slices 0-{k - 1} ({nb} bins), one slice per thread"}
        # the timed output against the CPU coder, byte for byte, on the first slice
        a, b = int(wl.tu_offsets[0]), int(wl.tu_offsets[1])
        pp = wl.params[0]
        eb = chk.encode_residual(wl.tus[a:b], wl.coeffs, int(pp["qp"]), int(pp["init_id"]), int(pp["flags"]))[0]
        e1 = int(d_entry[1])
        assert d_packed[:e1].cpu().numpy().tobytes() == eb.tobytes(), "slice 0: GPU bytes != CPU coder's"
        leg["verified"] += f"; slice 0 ({e1} bytes) == the CPU {kind} coder's"
    return leg


def _frame_batch(base, rows, frame_ids, dev):
    """Device-resident WPP batch made of the given frames of a base workload (frame f = substreams
    f * rows .. f * rows + rows - 1; its TUs and coefficients are contiguous in the base)."""
    import torch
    tu_off = base.tu_offsets.astype(np.int64)
    area = 1 << (base.tus["log2_w"].astype(np.int64) + base.tus["log2_h"].astype(np.int64))
    K = base.n_sub // rows
    f_tu = [(int(tu_off[f * rows]), int(tu_off[(f + 1) * rows])) for f in range(K)]
    f_co = []
    for a, b in f_tu:
        lo = int(base.tus["coeff_offset"][a])
        f_co.append((lo, int(base.tus["coeff_offset"][b - 1]) + int(area[b - 1])))
    tus_l, off_l, par_l, co_ranges = [], [0], [], []
    tu_base = co_base = 0
    for g in frame_ids:
        f = int(g) % K
        a, b = f_tu[f]
        lo, hi = f_co[f]
        t = base.tus[a:b].copy()
        t["coeff_offset"] = t["coeff_offset"] - np.uint64(lo) + np.uint64(co_base)
        tus_l.append(t)
        off_l.extend((tu_off[f * rows + 1:(f + 1) * rows + 1] - a + tu_base).tolist())
        par_l.append(base.params[f * rows:(f + 1) * rows])
        co_ranges.append((lo, hi))
        tu_base += b - a
        co_base += hi - lo
    d_base = torch.from_numpy(base.coeffs).to(dev)
    d_co = torch.cat([d_base[lo:hi] for lo, hi in co_ranges]) if co_ranges else d_base[:0].clone()
    del d_base
    return np.concatenate(tus_l), np.asarray(off_l, np.uint64), np.concatenate(par_l), d_co


def leg_wpp(args, r):
    """configs[3]: 4K WPP, one substream per CTU row with context hand-over after the first CTU of the row
    above (cabac_*_wpp: one launch), a 256-frame batch sharded frame-atomically over the ranks
    (entropy_coding_b200.sharding), entry points gathered on the host."""
    import torch
    from entropy_coding_b200 import sharding
    from entropy_coding_b200 import workload as W
    from entropy_coding_b200.types import WPP_DT
    from entropy_coding_b200.wpp import sync_records
    eng, dev, stream, world, rank = r["eng"], r["dev"], r["stream"], r["world"], r["rank"]
    K = 4  # distinct seeded frames; the job's frames cycle through them (generation time, not device time)
    base, rows, first = W.frames_as_ctu_rows(K, 3840, 2160, seed0=4000)
    total_frames = args.wpp_frames
    frame_cost = sharding.substream_costs(base.tus, base.tu_offsets).reshape(K, rows).sum(axis=1)
    bounds = sharding.partition_units(frame_cost[np.arange(total_frames) % K], world)  # a frame is the atomic unit
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    ids = np.arange(lo, hi)
    tus, tu_off, params, d_co = _frame_batch(base, rows, ids, dev)
    n_sub, n_tu, n_co = len(params), len(tus), int(d_co.numel())
    first_all = np.concatenate([first[(g % K) * rows:(g % K + 1) * rows] for g in ids]) if len(ids) else np.zeros(0, np.int64)
    wpp = sync_records(n_sub, rows, first_all)
    wl = W.Workload(tus, tu_off, params, np.zeros(0, np.int32))
    cap = wl.out_capacity()
    d_tus, d_tuoff, d_par, d_cap, d_wpp = (dev_bytes(dev, a) for a in (tus, tu_off, params, cap, wpp))
    d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
    d_packed = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
    d_entry = torch.zeros(n_sub + 1, dtype=torch.int64, device=dev)
    d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev)
    d_bins, d_st, d_dbins = torch.zeros_like(d_sizes), torch.zeros_like(d_sizes), torch.zeros_like(d_sizes)
    d_dec = torch.empty(n_co, dtype=torch.int32, device=dev)

    def enc():
        eng.encode_residual_wpp_dev(d_tus, d_tuoff, d_par, d_wpp, n_sub, d_co, d_out, d_cap, d_sizes, d_bins, status=d_st)
        eng.pack_dev(d_out, d_cap, d_sizes, n_sub, d_packed, d_entry)

    def dec():
        eng.decode_residual_wpp_dev(d_packed, d_entry, d_tus, d_tuoff, d_par, d_wpp, n_sub, d_dec, d_dbins, status=d_st)

    l0 = eng.launch_count
    t_enc = _time_passes(stream, enc, warm=1, reps=2)
    # the one cross-rank step of the path: per-substream byte counts -> global entry points (host side)
    t0 = time.perf_counter()
    sizes = d_sizes.cpu().numpy()
    entry_global, first_global = sharding.gather_entry_points(sizes)
    t_gather = time.perf_counter() - t0
    assert int(d_st.abs().max()) == 0, "wpp encode status"
    t_dec = _time_passes(stream, dec, warm=1, reps=2)
    launches = int(eng.launch_count - l0)
    assert int(d_st.abs().max()) == 0 and torch.equal(d_dec, d_co) and torch.equal(d_bins, d_dbins), "wpp leg round trip"
    bins = int(d_bins.to(torch.int64).sum())
    nbytes = int(sizes.astype(np.int64).sum())
    assert int(entry_global[-1]) >= nbytes and len(entry_global) == total_frames * rows + 1
    tb = torch.tensor([float(bins), float(nbytes)], dtype=torch.float64, device=dev)
    tm = torch.tensor([t_enc, t_dec, t_gather], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tb, op=dist.ReduceOp.SUM)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    bins_all, bytes_all = float(tb[0]), float(tb[1])
    t_enc_m, t_dec_m = float(tm[0]), float(tm[1])
    alg = 4 * n_co + 16 * n_tu + nbytes
    leg = {"config": f"BASELINE configs[3]: 4K (3840x2160) WPP, {rows} CTU rows per frame, {total_frames}-frame batch "
                     f"sharded over {world} GPU(s) frame-atomically; frames cycle through {K} seeded frames",
           "scaling": "strong", "frames_total": total_frames, "frames_this_rank": int(hi - lo),
           "substreams_total": total_frames * rows, "bins_per_pass_total": bins_all, "bitstream_bytes_total": bytes_all,
           "encode_mbins_s": bins_all / t_enc_m / 1e3, "decode_mbins_s": bins_all / t_dec_m / 1e3,
           "encode_ms": t_enc_m, "decode_ms": t_dec_m, "output_mb_s": bytes_all / t_enc_m / 1e3,
           "entry_point_gather_ms": 1e3 * float(tm[2]),
           "encode": _fractions(r, "encode_residual_kernel", bins, t_enc, alg),
           "decode": _fractions(r, "decode_residual_kernel", bins, t_dec, alg),
           "gpu_launches": launches,
           "api": "cabac_encode_residual_wpp_dev + cabac_pack_dev + cabac_decode_residual_wpp_dev (one coding launch "
                  "each; hand-over on the device); sharding.partition_units + sharding.gather_entry_points",
           "verified": "decode(encode(x)) == x, equal bin counts, global entry points cover every rank's bytes"}
    if rank == 0 and not args.no_cpu_baseline:
        # hand-over against the oracle run row after row, on the first frame of this rank
        chk, kind = cpu_checker()
        ent = d_entry[:rows + 1].cpu().numpy().astype(np.int64)
        got = d_packed[:int(ent[rows])].cpu().numpy()
        sync = None
        from entropy_coding_b200.types import SUB_ALIGN, SUB_CTX_IN, SUB_TERMINATE
        for rr in range(min(rows, 3)):
            a, b = int(tu_off[rr]), int(tu_off[rr + 1])
            pp = params[rr]
            fl = int(pp["flags"]) | (SUB_CTX_IN if sync is not None else 0)
            t_rows = tus[a:b].copy()
            lo_c = int(t_rows["coeff_offset"][0])
            hi_c = int(t_rows["coeff_offset"][-1]) + (1 << int(t_rows["log2_w"][-1] + t_rows["log2_h"][-1]))
            co = d_co[lo_c:hi_c].cpu().numpy()
            t_rows["coeff_offset"] -= np.uint64(lo_c)
            eb = chk.encode_residual(t_rows, co, int(pp["qp"]), int(pp["init_id"]), fl, ctx_in=sync)[0]
            assert got[ent[rr]:ent[rr + 1]].tobytes() == eb.tobytes(), f"WPP row {rr}: GPU bytes != CPU {kind}"
            k = int(first_all[rr])
            sync = chk.encode_residual(t_rows[:k], co, int(pp["qp"]), int(pp["init_id"]), fl & ~(SUB_TERMINATE | SUB_ALIGN),
                                       ctx_in=sync, want_ctx=True)[-1]
        leg["verified"] += f"; rows 0-{min(rows, 3) - 1} of frame 0 byte-exact vs the CPU {kind} coder run row after row"
        threads = host_threads()
        sample = W.Workload(tus[:int(tu_off[rows])], tu_off[:rows + 1], params[:rows], d_co[:int(base.coeffs.size // K)].cpu().numpy())
        cpu_pass(chk, sample, threads)
        te, td, nb = cpu_pass(chk, sample, threads)
        leg["cpu_baseline"] = {"value": 2 * nb / (te + td) / 1e6, "unit": UNIT, "cores": min(threads, rows), "kind": kind,
                               "encode_mbins_s": nb / te / 1e6, "decode_mbins_s": nb / td / 1e6,
                               "sample": f"the {rows} rows of frame 0 ({nb} bins) coded as independent rows, one per thread "
                                         "(the reference library has no wavefront scheduler; hand-over copies are not timed)"}
    del d_out, d_packed, d_dec, d_co
    torch.cuda.empty_cache()
    return leg


def leg_decode_sweep(args, r):
    """configs[4]: decode-only sweep over the substream count per launch (10^4, 10^5, 10^6), substreams of 64-256
    TUs, mixed QP presets.  The bitstreams of the seeded base batch (10^4 substreams) are produced once,
    checked against the CPU reference coder on a sample, and tiled on the device."""
    import torch
    from entropy_coding_b200 import workload as W
    eng, dev, stream = r["eng"], r["dev"], r["stream"]
    base = W.sweep_substreams(10_000, 42 + 10_000)
    cap1 = base.out_capacity()
    nco1 = int(base.coeffs.size)
    d_base = torch.from_numpy(base.coeffs).to(dev)
    rows = []
    chk = kind = None
    cpu = None
    if r["rank"] == 0 and not args.no_cpu_baseline:
        chk, kind = cpu_checker()
    for tiles in (1, 10, 100):
        n_sub = base.n_sub * tiles
        if n_sub > args.sweep_max:
            break
        tus, tu_off, params, cap = W.tile_descriptors(base, tiles, cap1)
        d_tus, d_tuoff, d_par, d_cap = (dev_bytes(dev, a) for a in (tus, tu_off, params, cap))
        n_tu = len(tus)
        del tus
        d_out = torch.empty(int(cap1[-1]), dtype=torch.uint8, device=dev)
        d_sizes1 = torch.zeros(base.n_sub, dtype=torch.int32, device=dev)
        d_bins1 = torch.zeros_like(d_sizes1)
        d_packed1 = torch.empty(int(cap1[-1]), dtype=torch.uint8, device=dev)
        d_entry1 = torch.zeros(base.n_sub + 1, dtype=torch.int64, device=dev)
        # bitstreams of the base batch (one tile), then replicated: tile k's substreams sit at k * bytes1 + entry1
        eng.encode_residual_dev(d_tus, d_tuoff, d_par, base.n_sub, d_base, d_out, d_cap, d_sizes1, d_bins1)
        eng.pack_dev(d_out, d_cap, d_sizes1, base.n_sub, d_packed1, d_entry1)
        torch.cuda.synchronize()
        bytes1 = int(d_entry1[-1])
        if chk is not None and tiles == 1:
            ent = d_entry1.cpu().numpy().astype(np.int64)
            pk = d_packed1[:bytes1].cpu().numpy()
            for i in range(0, base.n_sub, 250):
                a, b = int(base.tu_offsets[i]), int(base.tu_offsets[i + 1])
                pp = base.params[i]
                eb = chk.encode_residual(base.tus[a:b], base.coeffs, int(pp["qp"]), int(pp["init_id"]), int(pp["flags"]))[0]
                assert pk[ent[i]:ent[i + 1]].tobytes() == eb.tobytes(), f"sweep substream {i}: bytes != CPU {kind}"
            # CPU decode of a sample of the same bitstreams
            k = 1500
            sample = base.subset(range(k))
            out = np.zeros(int(ent[k]), np.uint8)
            out[:] = pk[:int(ent[k])]
            capk = ent[:k + 1].astype(np.uint64)
            szk = np.diff(ent[:k + 1]).astype(np.uint32)
            decoded = np.zeros(sample.coeffs.size, np.int32)
            threads = host_threads()
            chk.time_decode_residual(out, capk, szk, sample.tus, sample.tu_offsets, sample.params, decoded, threads)
            td = chk.time_decode_residual(out, capk, szk, sample.tus, sample.tu_offsets, sample.params, decoded, threads)
            assert (decoded == sample.coeffs).all()
            nbk = int(d_bins1[:k].to(torch.int64).sum())
            cpu = {"value": nbk / td / 1e6, "unit": UNIT, "cores": threads, "kind": kind,
                   "sample": f"substreams 0-{k - 1} of the base batch ({nbk} bins), decode only, {threads} threads, "
                             "bitstreams byte-identical to the CPU coder's (every 250th substream compared)"}
        d_in = d_packed1[:bytes1].repeat(tiles)
        ent1 = d_entry1[:-1]
        d_entry = torch.cat([(ent1[None, :] + (torch.arange(tiles, device=dev, dtype=torch.int64) * bytes1)[:, None]).reshape(-1),
                             torch.tensor([bytes1 * tiles], device=dev, dtype=torch.int64)])
        del d_out, d_packed1
        d_dec = torch.empty(nco1 * tiles, dtype=torch.int32, device=dev)
        d_dbins = torch.zeros(n_sub, dtype=torch.int32, device=dev)
        d_st = torch.zeros(n_sub, dtype=torch.int32, device=dev)

        def dec():
            eng.decode_residual_dev(d_in, d_entry, d_tus, d_tuoff, d_par, n_sub, d_dec, d_dbins, status=d_st)

        t_dec = _time_passes(stream, dec, warm=1, reps=2)
        assert int(d_st.abs().max()) == 0, "sweep decode status"
        assert bool((d_dec.view(tiles, nco1) == d_base[None, :]).all()), "sweep decode != the encoder's input"
        bins = int(d_dbins.to(torch.int64).sum())
        alg = 4 * nco1 * tiles + 16 * n_tu + bytes1 * tiles
        row = {"substreams": n_sub, "tus": n_tu, "bins": bins, "bitstream_bytes": bytes1 * tiles,
               "decode_ms": t_dec, "decode_mbins_s": bins * r["world"] / t_dec / 1e3,
               "hbm_gbs": alg / (t_dec * 1e-3) / 1e9}
        row.update(_fractions(r, "decode_residual_kernel", bins, t_dec, alg))
        rows.append(row)
        del d_in, d_entry, d_dec, d_dbins, d_st, d_tus, d_tuoff, d_par, d_cap
        torch.cuda.empty_cache()
    leg = {"config": "BASELINE configs[4]: decode-only sweep, independent substreams of 64-256 TUs (4x4 .. 16x16, luma + "
                     "chroma), mixed qp0/qp32/qp51 presets; a seeded base batch of 10^4 substreams tiled on the device",
           "per_gpu": True, "points": rows,
           "verified": "every decoded tile == the encoder's input; status clear"}
    if cpu:
        leg["cpu_baseline"] = cpu
    del d_base
    torch.cuda.empty_cache()
    return leg


def leg_slice_decode(args, r):
    """whole slices decoded from their bytes alone (cabac_decode_slices_dev: CTU-level syntax
    + residuals on the device).  Input: the VTM-produced fixture slices of tests/golden/vtm_*.npz (real slice payloads
    of the reference's encoder; QCIF and 416x240), replicated on the device to 64 / 8192 substreams per launch - the
    same bytes many times over, so the figure is a kernel rate, not a statement about a particular video.  Every copy's
    CU / TU counts, bin count and coefficients are compared with what the reference's writer recorded."""
    import torch
    import entropy_coding_b200 as E
    from entropy_coding_b200.types import CTX_DT, PARAMS_DT, SL_WPP, SLICE_DT, SUB_CTX_IN, SUB_TERMINATE
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from vtm_streams import load_fixture
    import glob
    eng, stream, dev = r["eng"], r["stream"], r["dev"]
    items = []
    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "vtm_*.npz"))):
        for s in load_fixture(os.path.basename(path))[0]:
            if "slice" not in s or not len(s["slice"]):
                continue
            sp = s["slice"].astype(np.uint32).copy()
            if int(sp[30]) & SL_WPP:
                if int(sp[3]) != 0:
                    continue
                sp[30] &= ~np.uint32(SL_WPP)
            items.append((s, sp.view(SLICE_DT)[0]))
    if not items:
        return {"unavailable": "no slice fixtures under tests/golden"}
    rows = []
    for n in (64, 8192):
        order = [k % len(items) for k in range(n)]
        params, sps, ctx_in = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT), np.zeros((n, 379), CTX_DT)
        data, off = [], [0]
        for j, k in enumerate(order):
            s, sp = items[k]
            params[j] = (s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)
            sps[j] = sp
            ctx_in[j]["state0"], ctx_in[j]["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
            data.append(s["payload"])
            off.append(off[-1] + s["payload"].size)
        cap = E.CabacEngine.slice_capacities(sps)
        d = {k: dev_bytes(dev, v) for k, v in (("in", np.concatenate(data)), ("off", np.asarray(off, np.uint64)),
                                               ("p", params), ("sp", sps), ("ci", ctx_in), ("cu", cap["cu"]),
                                               ("tu", cap["tu"]), ("ctu", cap["ctu"]), ("co", cap["coeff"]))}
        cus = torch.empty(int(cap["cu"][-1]) * 32, dtype=torch.uint8, device=dev)
        tus = torch.empty(int(cap["tu"][-1]) * 16, dtype=torch.uint8, device=dev)
        tuf = torch.empty(int(cap["tu"][-1]), dtype=torch.uint8, device=dev)
        ctus = torch.empty(int(cap["ctu"][-1]) * 32, dtype=torch.uint8, device=dev)
        co = torch.zeros(int(cap["coeff"][-1]), dtype=torch.int32, device=dev)
        cnt = torch.zeros(n * 3, dtype=torch.int32, device=dev)
        nb = torch.zeros(n, dtype=torch.int32, device=dev)
        st = torch.zeros(n, dtype=torch.int32, device=dev)

        def dec():
            eng.decode_slices_dev(d["in"], d["off"], d["p"], d["sp"], n, d["cu"], d["tu"], d["ctu"], d["co"], cus, tus, tuf,
                                  ctus, co, counts=cnt, num_bins=nb, ctx_in=d["ci"], status=st)

        t = _time_passes(stream, dec, warm=2, reps=3)
        assert int(st.abs().max()) == 0, "slice decode status"
        counts, bins_h, co_h = cnt.cpu().numpy().reshape(n, 3), nb.cpu().numpy(), co.cpu().numpy()
        n_coeff = 0
        for j, k in enumerate(order):
            s, _ = items[k]
            assert int(bins_h[j]) == s["num_bins"] and int(counts[j, 0]) == len(s["cus"]) and int(counts[j, 1]) == len(s["tus"])
            k0 = int(cap["coeff"][j])
            assert (co_h[k0:k0 + s["coeffs"].size] == s["coeffs"]).all(), f"slice copy {j}: coefficients"
            n_coeff += s["coeffs"].size
        bins = int(bins_h.astype(np.int64).sum())
        alg = off[-1] + 4 * n_coeff + 16 * int(counts[:, 1].sum()) + 32 * int(counts[:, 0].sum())
        row = {"substreams": n, "bins": bins, "bitstream_bytes": off[-1], "cus": int(counts[:, 0].sum()),
               "tus": int(counts[:, 1].sum()), "coefficients_out": n_coeff, "decode_ms": t,
               "decode_mbins_s": bins * r["world"] / t / 1e3, "hbm_gbs": alg / (t * 1e-3) / 1e9}
        row["hbm_frac"] = row["hbm_gbs"] / load_peaks()[0]
        # the writer direction on the same records (cabac_encode_slices_dev): records -> the writer's bytes
        from entropy_coding_b200.types import CU_DT, SUB_ALIGN, TU_DT
        cus_h, tus_h = cus.cpu().numpy().view(CU_DT), tus.cpu().numpy().view(TU_DT)
        tuf_h = tuf.cpu().numpy()
        pick = lambda a, o, col: np.concatenate([a[int(o[j]):int(o[j]) + int(counts[j, col])] for j in range(n)])  # noqa: E731
        exact = lambda col: np.concatenate([[0], np.cumsum(counts[:, col].astype(np.int64))]).astype(np.uint64)  # noqa: E731
        wp = params.copy()
        wp["flags"] |= SUB_ALIGN
        out_off = np.concatenate([[0], np.cumsum([(p.size + 79) // 16 * 16 for p in data])]).astype(np.uint64)
        words = int(bins * 1.5) + 64 * n
        w = {k: dev_bytes(dev, v) for k, v in (("p", wp), ("cus", pick(cus_h, cap["cu"], 0)), ("cuo", exact(0)),
                                               ("tus", pick(tus_h, cap["tu"], 1)), ("tuf", pick(tuf_h, cap["tu"], 1)),
                                               ("tuo", exact(1)), ("oo", out_off))}
        cmd = torch.empty(words, dtype=torch.int32, device=dev)
        out = torch.zeros(int(out_off[-1]), dtype=torch.uint8, device=dev)
        osz = torch.zeros(n, dtype=torch.int32, device=dev)
        wst = torch.zeros(n, dtype=torch.int32, device=dev)
        wnb = torch.zeros(n, dtype=torch.int32, device=dev)

        def enc():
            eng.encode_slices_dev(w["p"], d["sp"], n, w["cus"], w["cuo"], w["tus"], w["tuf"], w["tuo"], ctus, d["ctu"], co, cmd,
                                  words, out, w["oo"], osz, wst, num_bins=wnb, ctx_in=d["ci"])

        te = _time_passes(stream, enc, warm=2, reps=3)
        assert int(wst.abs().max()) == 0, "slice encode status"
        out_h, osz_h = out.cpu().numpy(), osz.cpu().numpy()
        for j, k in enumerate(order):
            a = int(out_off[j])
            assert out_h[a:a + int(osz_h[j])].tobytes() == items[k][0]["payload"].tobytes(), f"slice copy {j}: bytes"
        assert (wnb.cpu().numpy() == bins_h).all()
        row["encode_ms"], row["encode_mbins_s"] = te, bins * r["world"] / te / 1e3
        rows.append(row)
        del cus, tus, tuf, ctus, co, cnt, nb, st, d, w, cmd, out, osz, wst, wnb
        torch.cuda.empty_cache()
    rows_gen = _slice_decode_generated(r)
    cpu = None
    if r["rank"] == 0 and not args.no_cpu_baseline:
        # picture-sized slices keep the per-call Python overhead out of the CPU figure
        gen_path = os.path.join(ROOT, "tests", "golden", "gen_slices.npz")
        cpu_items = items
        if os.path.exists(gen_path):
            z = np.load(gen_path)
            cpu_items = []
            for k in range(2):
                meta = z[f"hd{k}_meta"]
                cpu_items.append((dict(payload=z[f"hd{k}_payload"], qp=int(meta[0]), init_id=int(meta[1]), flags=int(meta[2]),
                                       num_bins=int(meta[3]), ctx_in=None), z[f"hd{k}_slice"].view(SLICE_DT)[0]))
        cpu = _cpu_slice_decode(cpu_items)
    leg = _slice_leg(rows, items)
    if rows_gen:
        leg["generated"] = {"input": "tests/golden/gen_slices.npz (+ gen_slices_wpp.npz: the 1080p pictures with WPP, nine row substreams per slice, "
                                     "written through cabac_encode_slices_wpp_dev): legal all-intra slices from the stream generator "
                                     "(tests/golden/make_gen_slices.py), dense residuals (1.5 bins per coefficient "
                                     "position), replicated on the device; ground truth = the generating walk's records",
                            "points": rows_gen,
                            "verified": "every copy: bin, CU and TU counts; slice 0 of every set: coefficient checksum; encode_* = "
                                    "cabac_encode_slices_dev on the decoded records: every copy's bytes == the generated payload"}
    if cpu:
        leg["cpu_baseline"] = cpu
    return leg


def _slice_decode_generated(r):
    """cabac_decode_slices on the committed generated slices: 64 x 1080p pictures as one slice each (the shape of
    BASELINE configs[2], decoded from the bytes alone) and 2 048 x 416x240."""
    import torch
    import entropy_coding_b200 as E
    from entropy_coding_b200.types import PARAMS_DT, SLICE_DT, SUB_TERMINATE
    path = os.path.join(ROOT, "tests", "golden", "gen_slices.npz")
    if not os.path.exists(path):
        return None
    z = dict(np.load(path))
    sets = list(z["sets"])
    legs = [("hd", 64), ("wqvga", 2048)]
    wpath = os.path.join(ROOT, "tests", "golden", "gen_slices_wpp.npz")
    if os.path.exists(wpath):  # the 1080p pictures again with entropy_coding_sync_enabled: nine row substreams each
        zw = np.load(wpath)
        z.update({k: zw[k] for k in zw.files if k != "sets"})
        sets += list(zw["sets"])
        legs.append(("hdwpp", 64))
    eng, stream, dev = r["eng"], r["stream"], r["dev"]
    rows = []
    for entry, n in legs:
        cnt_set = next(int(t.split(":")[1]) for t in sets if t.startswith(entry + ":"))
        base = [(z[f"{entry}{k}_payload"], z[f"{entry}{k}_slice"].view(SLICE_DT)[0], z[f"{entry}{k}_meta"],
                 int(z[f"{entry}{k}_checksum"])) for k in range(cnt_set)]
        params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
        data, off = [], [0]
        for j in range(n):
            pl, sp, meta, _ = base[j % len(base)]
            params[j] = (int(meta[0]), int(meta[1]), int(meta[2]) | SUB_TERMINATE, 0)
            sps[j] = sp
            data.append(pl)
            off.append(off[-1] + pl.size)
        cap = E.CabacEngine.slice_capacities(sps)
        d = {k: dev_bytes(dev, v) for k, v in (("in", np.concatenate(data)), ("off", np.asarray(off, np.uint64)), ("p", params),
                                               ("sp", sps), ("cu", cap["cu"]), ("tu", cap["tu"]), ("ctu", cap["ctu"]),
                                               ("co", cap["coeff"]))}
        cus = torch.empty(int(cap["cu"][-1]) * 32, dtype=torch.uint8, device=dev)
        tus = torch.empty(int(cap["tu"][-1]) * 16, dtype=torch.uint8, device=dev)
        tuf = torch.empty(int(cap["tu"][-1]), dtype=torch.uint8, device=dev)
        ctus = torch.empty(int(cap["ctu"][-1]) * 32, dtype=torch.uint8, device=dev)
        co = torch.zeros(int(cap["coeff"][-1]), dtype=torch.int32, device=dev)
        cnt = torch.zeros(n * 3, dtype=torch.int32, device=dev)
        nb = torch.zeros(n, dtype=torch.int32, device=dev)
        st = torch.zeros(n, dtype=torch.int32, device=dev)

        def dec():
            eng.decode_slices_dev(d["in"], d["off"], d["p"], d["sp"], n, d["cu"], d["tu"], d["ctu"], d["co"], cus, tus, tuf,
                                  ctus, co, counts=cnt, num_bins=nb, status=st)

        t = _time_passes(stream, dec, warm=1, reps=2)
        assert int(st.abs().max()) == 0, "generated slices: status"
        counts, bins_h = cnt.cpu().numpy().reshape(n, 3), nb.cpu().numpy()
        for j in range(n):
            meta = base[j % len(base)][2]
            assert int(bins_h[j]) == int(meta[3]) and int(counts[j, 0]) == int(meta[4]) and int(counts[j, 1]) == int(meta[5])
        k0, n0 = int(cap["coeff"][0]), int(base[0][2][6])
        c0 = co[k0:k0 + n0].to(torch.int64)
        chk = int((c0 * torch.arange(1, n0 + 1, device=dev, dtype=torch.int64)).sum()) & 0xFFFFFFFFFFFFFFFF
        assert chk == base[0][3], "generated slices: coefficient checksum"
        bins = int(bins_h.astype(np.int64).sum())
        ncoef = sum(int(base[j % len(base)][2][6]) for j in range(n))
        alg = off[-1] + 4 * ncoef + 16 * int(counts[:, 1].sum()) + 32 * int(counts[:, 0].sum())
        # the writer direction on the records just decoded (compacted on the device): bytes == the generated payloads
        from entropy_coding_b200.types import SUB_ALIGN
        cu_cap, tu_cap = cap["cu"].astype(np.int64), cap["tu"].astype(np.int64)
        cus2 = torch.cat([cus.view(-1, 32)[int(cu_cap[j]):int(cu_cap[j]) + int(counts[j, 0])] for j in range(n)]).contiguous()
        tus2 = torch.cat([tus.view(-1, 16)[int(tu_cap[j]):int(tu_cap[j]) + int(counts[j, 1])] for j in range(n)]).contiguous()
        tuf2 = torch.cat([tuf[int(tu_cap[j]):int(tu_cap[j]) + int(counts[j, 1])] for j in range(n)]).contiguous()
        exact = lambda col: np.concatenate([[0], np.cumsum(counts[:, col].astype(np.int64))]).astype(np.uint64)  # noqa: E731
        wp = params.copy()
        wp["flags"] |= SUB_ALIGN
        words = bins + 64 * n
        w = {k: dev_bytes(dev, v) for k, v in (("p", wp), ("cuo", exact(0)), ("tuo", exact(1)))}
        cmd = torch.empty(words, dtype=torch.int32, device=dev)
        # row substreams: one per slice, or one per CTU row of a WPP slice (each row gets the slice's capacity)
        row_off = E.CabacEngine.slice_rows(sps)
        n_rows = int(row_off[-1])
        per_row = np.concatenate([[(data[j].size + 79) // 16 * 16] * int(row_off[j + 1] - row_off[j]) for j in range(n)])
        out_off = np.concatenate([[0], np.cumsum(per_row)]).astype(np.uint64)
        w["oo"], w["ro"] = dev_bytes(dev, out_off), dev_bytes(dev, row_off)
        out = torch.zeros(int(out_off[-1]), dtype=torch.uint8, device=dev)
        osz = torch.zeros(n_rows, dtype=torch.int32, device=dev)
        wst = torch.zeros(n_rows, dtype=torch.int32, device=dev)

        def enc():
            eng.encode_slices_wpp_dev(w["p"], d["sp"], n, cus2, w["cuo"], tus2, tuf2, w["tuo"], ctus, d["ctu"], co, cmd, words,
                                      w["ro"], n_rows, out, w["oo"], osz, wst)

        te = _time_passes(stream, enc, warm=1, reps=2)
        assert int(wst.abs().max()) == 0, "generated slices: writer status"
        out_h, osz_h = out.cpu().numpy(), osz.cpu().numpy()
        for j in range(n):
            got = np.concatenate([out_h[int(out_off[q]):int(out_off[q]) + int(osz_h[q])]
                                  for q in range(int(row_off[j]), int(row_off[j + 1]))])
            assert got.tobytes() == data[j].tobytes(), f"generated slice copy {j}: written bytes"
        del cus2, tus2, tuf2, cmd, out, osz, wst, w
        rows.append({"set": f"{n} x {int(sps[0]['pic_w'])}x{int(sps[0]['pic_h'])}" + (" WPP" if entry.endswith("wpp") else ""),
                     "substreams": n, "row_substreams": n_rows, "bins": bins, "encode_ms": te, "encode_mbins_s": bins * r["world"] / te / 1e3,
                     "bitstream_bytes": off[-1], "coefficients_out": ncoef, "decode_ms": t,
                     "decode_mbins_s": bins * r["world"] / t / 1e3, "hbm_gbs": alg / (t * 1e-3) / 1e9,
                     "hbm_frac": alg / (t * 1e-3) / 1e9 / load_peaks()[0],
                     "cycles_per_bin_per_warp": (t * 1e-3) * _sm_mhz(r) * 1e6 / (bins / n) if n <= 148 else None})
        del cus, tus, tuf, ctus, co, cnt, nb, st, d
        torch.cuda.empty_cache()
    return rows


def _cpu_slice_decode(items):
    """CPU figure beside the slice decoder: the same slices through oracle/libsyntax_host.so (the CTU-level walk on
    the oracle's scalar arithmetic / residual decoder), every host thread decoding copies for about ten seconds.
    kind "port": the reference's own CABACReader::coding_tree_unit needs VTM's CodingStructure / Picture around it
    and is not drivable from flat inputs."""
    import ctypes as C
    import time
    from concurrent.futures import ThreadPoolExecutor
    path = os.path.join(ROOT, "oracle", "libsyntax_host.so")
    if not os.path.exists(path):
        return None
    from entropy_coding_b200.types import CTU_DT, CTX_DT, CU_DT, PARAMS_DT, SLICE_DT, SUB_CTX_IN, SUB_TERMINATE, TU_DT
    lib = C.CDLL(path)
    lib.syn_decode_slice.restype = C.c_long
    threads = host_threads()

    def p(a):
        return a.ctypes.data_as(C.c_void_p)

    def worker(tid, deadline):
        bins = 0
        bufs = {}
        while time.perf_counter() < deadline:
            for k, (s, sp) in enumerate(items):
                if k not in bufs:
                    area = int(sp["x1"] - sp["x0"]) * int(sp["y1"] - sp["y0"])
                    cap = area // 16 * 2 + 8
                    ci = np.zeros(379, CTX_DT)
                    fl = s["flags"] | SUB_TERMINATE
                    if s["ctx_in"] is not None:
                        ci["state0"], ci["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
                        fl |= SUB_CTX_IN
                    bufs[k] = (np.array([(s["qp"], s["init_id"], fl, 0)], PARAMS_DT),
                               np.array([sp], SLICE_DT), ci, np.zeros(cap, CU_DT), np.zeros(cap, TU_DT),
                               np.zeros(cap, np.uint8), np.zeros(4096, CTU_DT), np.zeros(area * 3 // 2 + 64, np.int32),
                               np.zeros(3, np.uint32), np.zeros(1, np.uint32), np.zeros(379, CTX_DT),
                               np.zeros(1, np.uint32), np.zeros(1 << 16, np.uint32), np.ascontiguousarray(s["payload"]), cap)
                par, spa, ci, cus, tus, tuf, ctus, co, cnt, nb, cout, st, tr, pay, cap = bufs[k]
                nt = C.c_long(0)
                lib.syn_decode_slice(p(pay), C.c_long(pay.size), p(par), p(spa), p(ci), p(cus), C.c_long(cap), p(tus), p(tuf),
                                     C.c_long(cap), p(ctus), C.c_long(4096), p(co), C.c_long(co.size), p(cnt), p(nb), p(cout),
                                     p(st), p(tr), C.c_long(tr.size), C.byref(nt))
                assert int(st[0]) == 0 and int(nb[0]) == s["num_bins"]
                bins += int(nb[0])
        return bins

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        done = list(ex.map(lambda t: worker(t, t0 + 8.0), range(threads)))
    dt = time.perf_counter() - t0
    return {"value": sum(done) / dt / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{len(items)} slices ({sum(s['num_bins'] for s, _ in items)} bins) decoded over and over for {dt:.1f} s on {threads} threads "
                      "(oracle/libsyntax_host.so: the CTU-level walk on the oracle's scalar decoder)"}


def _slice_leg(rows, items):
    return {"config": "SURVEY section 8 f2 (the decode side of configs[0-1]): all-intra dual-tree slices decoded from the "
                      "bytes alone; VTM-produced fixture slices replicated on the device",
            "per_gpu": True, "points": rows,
            "api": "cabac_decode_slices_dev (decode_slice_kernel: one warp per slice); encode_*: the writer direction, "
                   "cabac_encode_slices_dev on the records the decode produced (emit_slice_kernel x 2 + the stream encoder)",
            "fixture_slices": len(items),
            "note": "64 slices per launch is the latency floor of one dependent chain per slice; 8192 fills the GPU. The "
                    "fixture slices are sparse (0.1-0.3 bins per coefficient position, large blocks with few levels), so "
                    "per-block work - zeroing and writing back whole blocks - weighs more here than in the headline batch",
            "verified": "every copy: bin count, CU and TU counts and all coefficients == the reference writer's records; "
                        "encode: every copy's bytes == the VTM payload"}


def run_e2e(args, r, i16=False):
    """The same step through the host-pointer C ABI: pinned host buffers in, pinned host buffers out.
    i16: the coefficient slabs cross the ABI as int16 (cabac_*_i16) instead of the reference's int32 TCoeff."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    eng, wl, cap = r["eng"], r["wl"], r["cap"]
    n_sub, n_tu, n_co = r["n_sub"], r["n_tu"], r["n_co"]
    lib, h = eng.lib, eng.h
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    h_tus, h_tuoff = eng.pinned_copy(wl.tus), eng.pinned_copy(wl.tu_offsets)
    h_par, h_cap = eng.pinned_copy(wl.params), eng.pinned_copy(cap)
    h_co = eng.pinned_copy(wl.coeffs.astype(np.int16) if i16 else wl.coeffs)
    f_enc = lib.cabac_encode_residual_packed_i16 if i16 else lib.cabac_encode_residual_packed
    f_dec = lib.cabac_decode_residual_i16 if i16 else lib.cabac_decode_residual
    packed_cap = int(r["out_bytes"] * 1.25) + 4096
    h_packed = eng.host_array(packed_cap, np.uint8)
    h_entry = eng.host_array(n_sub + 1, np.uint64)
    h_sizes, h_ebins, h_est = (eng.host_array(n_sub, np.uint32) for _ in range(3))
    h_eflags, h_dflags = eng.host_array(n_tu, np.uint8), eng.host_array(n_tu, np.uint8)
    h_dec = eng.host_array(n_co, np.int16 if i16 else np.int32)
    h_dbins, h_dbits, h_dst = (eng.host_array(n_sub, np.uint32) for _ in range(3))

    def step():
        # the batch description goes to the device once per step; both calls below reuse it
        rc = lib.cabac_pin_descriptors(h, P(h_tus), P(h_tuoff), P(h_par), n_sub)
        assert rc == 0, (rc, lib.cabac_last_error(h))
        rc = f_enc(h, P(h_tus), P(h_tuoff), P(h_par), n_sub, P(h_co), n_co, None,
                   P(h_cap), P(h_packed), packed_cap, P(h_entry), P(h_sizes),
                   P(h_ebins), P(h_eflags), None, P(h_est))
        assert rc == 0, (rc, lib.cabac_last_error(h))
        rc = f_dec(h, P(h_packed), P(h_entry), P(h_tus), P(h_tuoff), P(h_par), n_sub, None,
                   P(h_dec), n_co, P(h_dbins), P(h_dbits), P(h_dflags), None, P(h_dst))
        assert rc == 0, (rc, lib.cabac_last_error(h))

    for _ in range(max(args.warmup, 3)):
        step()
    r["barrier"]()
    launches0 = eng.launch_count
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    r["barrier"]()
    lib.cabac_unpin_descriptors(h)
    assert (h_dec.astype(np.int32) == wl.coeffs).all() and (h_ebins == h_dbins).all(), "e2e round trip failed"
    total = int(h_entry[n_sub])
    tt = torch.tensor([dt], dtype=torch.float64, device=r["dev"])
    if r["world"] > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    small = h_tus.nbytes + h_tuoff.nbytes + h_par.nbytes
    h2d = (h_co.nbytes + small + h_cap.nbytes) + (total + h_entry.nbytes)  # descriptors once per step (cabac_pin_descriptors)
    d2h = (total + h_entry.nbytes + 3 * h_sizes.nbytes + h_eflags.nbytes) + (h_dec.nbytes + 3 * h_dbins.nbytes + h_dflags.nbytes)
    return {"value": r["all_bins"] * args.steps / float(tt) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * float(tt) / args.steps,
            "gpu_launches": int(eng.launch_count - launches0),
            "coefficient_type": "int16" if i16 else "int32 (TCoeff)",
            "api": ("cabac_pin_descriptors + cabac_encode_residual_packed%s + cabac_decode_residual%s (host pointers; the batch "
                    "description is uploaded once per step, pinned coefficient slabs are read / written by the kernels "
                    "directly over PCIe)") % (("_i16", "_i16") if i16 else ("", ""))}


def run_cpu_baseline(args, r):
    chk, kind = cpu_checker()
    threads = host_threads()
    frames = max(1, min(args.cpu_sample_frames, args.frames))
    per_frame = r["n_sub"] // args.frames
    sample = r["wl"].subset(range(frames * per_frame))
    cpu_pass(chk, sample, threads)  # warm-up (first-touch pages, thread start-up)
    t_enc, t_dec, bins = cpu_pass(chk, sample, threads)
    t1_enc, t1_dec, bins1 = cpu_pass(chk, r["wl"].subset(range(per_frame)), 1)
    return {"value": 2 * bins / (t_enc + t_dec) / 1e6, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"frames 0-{frames - 1} of the timed batch ({sample.n_sub} substreams, {bins} bins), "
                      f"encode+decode, {threads} threads, dynamic substream hand-out",
            "encode_mbins_s": bins / t_enc / 1e6, "decode_mbins_s": bins / t_dec / 1e6,
            "single_thread_mbins_s": 2 * bins1 / (t1_enc + t1_dec) / 1e6}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    r = run_b200_arm(args)
    e2e = None if args.no_e2e else run_e2e(args, r)
    e2e_i16 = None if args.no_e2e else run_e2e(args, r, i16=True)
    legs = {}
    if not args.no_legs and args.substreams == "ctu":
        legs["1080p_64_slices"] = r.get("leg_slices")
        legs["4k_wpp"] = leg_wpp(args, r)
        legs["decode_sweep"] = leg_decode_sweep(args, r)
        legs["slice_decode"] = leg_slice_decode(args, r)
    if r["rank"] != 0:
        return
    peak, peak_src = load_peaks()
    facts = load_profile_facts()
    tk = r["t_kernel"]
    dom = max(("encode_residual_kernel", "decode_residual_kernel"), key=lambda k: tk[k])
    # algorithmic bytes per launch: int32 coefficient positions + 16 B per TU
    # descriptor + bitstream bytes (written by the encoder / read by the decoder)
    alg = 4 * r["n_co"] + 16 * r["n_tu"] + r["out_bytes"]
    achieved = alg / (tk[dom] * 1e-3) / 1e9
    sm_mhz = (r["clk"].get("sm_mhz") or 1965.0)
    roof = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "traffic": facts.get(dom, {}).get("dram_bytes_per_launch"),
            "peak_source": peak_src, "algorithmic_bytes_per_launch": alg,
            "kernel_ms": {k: round(v, 4) for k, v in tk.items()},
            "note": "integer/latency-bound serial recurrence: see issue_roofline; HBM is not the binding resource"}
    ipb = facts.get(dom, {}).get("warp_instr_per_bin")
    issue = {"kernel": dom, "warp_instr_per_bin": ipb, "sm_count": 148, "sm_mhz": sm_mhz,
             "mbins_s_achieved": r["enc_bins"] / (tk[dom] * 1e-3) / 1e6}
    if ipb:
        # 4 warp-instruction issue slots per SM per clock
        bound = 148 * 4 * sm_mhz * 1e6 / ipb / 1e6
        issue.update({"mbins_s_bound": bound, "frac": issue["mbins_s_achieved"] / bound})
    per_kernel = {}
    for k in ("encode_residual_kernel", "decode_residual_kernel"):
        f = facts.get(k, {})
        if f.get("warp_instr_per_bin"):
            b = 148 * 4 * sm_mhz * 1e6 / f["warp_instr_per_bin"] / 1e6
            a = r["enc_bins"] / (tk[k] * 1e-3) / 1e6
            per_kernel[k] = {"warp_instr_per_bin": f["warp_instr_per_bin"], "mbins_s_achieved": a, "mbins_s_bound": b,
                             "frac": a / b, "issue_active_pct_ncu": f.get("issue_active_pct"),
                             "shared_bank_conflicts_ncu": f.get("shared_bank_conflicts"),
                             "dram_bytes_ncu": f.get("dram_bytes_per_launch")}
    for k, v in per_kernel.items():
        tpw = facts.get(k, {}).get("thread_inst_per_warp_inst")
        if tpw:
            tib = v["warp_instr_per_bin"] * tpw
            lane_bound = 148 * 128 * sm_mhz * 1e6 / tib / 1e6
            v.update({"thread_inst_per_warp_inst": tpw, "thread_instr_per_bin": tib, "lane_mbins_s_bound": lane_bound,
                      "lane_frac": v["mbins_s_achieved"] / lane_bound})
    if dom in per_kernel and "lane_frac" in per_kernel[dom]:
        issue.update({"lane_frac": per_kernel[dom]["lane_frac"], "lane_mbins_s_bound": per_kernel[dom]["lane_mbins_s_bound"],
                      "thread_instr_per_bin": per_kernel[dom]["thread_instr_per_bin"]})
    issue["definitions"] = ("frac = achieved / (148 SMs x 4 warp-issue slots x clock / warp-instructions per bin); lane_frac = "
                            "achieved / (148 x 128 lanes x clock / thread-instructions per bin): the first says how busy the "
                            "issue ports are with THIS instruction stream, the second counts the lanes as well - it equals "
                            "frac because the lockstep kernels run 25-31 of 32 lanes on every instruction, most of them "
                            "repeating the same serial arithmetic (profiles/r2_decoder_modes.md: one thread per substream "
                            "needs 4.8x fewer warp instructions but runs 7.4 lanes wide)")
    issue["kernels"] = per_kernel
    issue["source"] = "instructions per bin, issue activity, bank conflicts and DRAM bytes: profiles/kernel_facts.json (ncu)"
    line = {
        "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": r["world"], "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": WORKLOAD if args.substreams == "ctu" else WORKLOAD.replace(
                       "one substream per 128x128 CTU (8640 substreams/GPU)", "one substream per frame"),
                   "frames_per_gpu": args.frames, "substreams_per_gpu": r["n_sub"],
                   "tus_per_gpu": r["n_tu"], "coeff_positions_per_gpu": r["n_co"],
                   "bins_per_step_per_gpu": 2 * r["enc_bins"], "bitstream_bytes_per_gpu": r["out_bytes"],
                   "l2": "inputs larger than L2 (coefficient slab %.0f MB per pass)" % (4 * r["n_co"] / 1e6),
                   "step": "encode_residual + pack + decode_residual",
                   "sample": "the whole batch every step (the reference arm codes a bounded sample of the same batch per "
                             "step - its `config.sample` names the frames; both lines are bins per second, so the ratio "
                             "compares like with like)"},
        "encode_mbins_s": r["enc_bins"] / (tk["encode_residual_kernel"] + tk["pack"]) / 1e3,
        "decode_mbins_s": r["enc_bins"] / tk["decode_residual_kernel"] / 1e3,
        "output_mb_s": r["out_bytes"] / (tk["encode_residual_kernel"] + tk["pack"]) / 1e3,
        "clocks": r["clk"], "gpu_launches": int(r["launches"]), "roofline": roof, "issue_roofline": issue,
        "verified": r["verified"],
        "host_placement": r.get("numa", "n/a"),
        "decode_scope": "the headline's decode half is cabac_decode_residual: substreams made of TUs only (the TU list is an "
                        "input). A real VVC slice interleaves CTU-level syntax on the same arithmetic-decoder state: that "
                        "path is cabac_decode_slices (SURVEY section 8 row f2, all-intra dual-tree slices), measured in "
                        "configs.slice_decode on VTM-produced slices; encode_mbins_s drops into a codec as it is",
    }
    if legs:
        line["configs"] = legs
    if e2e:
        line["e2e"] = e2e
        line["e2e"]["limiter"] = ("host side: the step's 1.85 GB per rank cross PCIe into ONE host memory system; plain pinned "
                                  "cudaMemcpyAsync of the same bytes on 1/2/4/8 ranks takes 21.8 / 38.0 / 74.8 / 120.0 ms "
                                  "(profiles/r2_pcie_ceiling.txt) - from four ranks up the end-to-end step runs at that ceiling")
    if e2e_i16:
        line["e2e_i16"] = e2e_i16
    if r["world"] == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = run_cpu_baseline(args, r)
    if _SAVED_STDOUT is not None:
        sys.stdout.flush()
        os.dup2(_SAVED_STDOUT, 1)
    print(json.dumps(line), flush=True)


def _shutdown():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    try:
        main()
    finally:
        _shutdown()
