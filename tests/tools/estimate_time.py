"""Dev tool: device time of the bit estimator next to the encoder on the bench batch, and on an
RDO-shaped batch (many short candidate TU lists priced from saved contexts)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

dev = torch.device("cuda", 0)
eng = E.CabacEngine(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); eng.set_stream(st.cuda_stream)
B = lambda a: torch.from_numpy(np.frombuffer(a.tobytes(), np.uint8).copy()).to(dev)
ev = lambda: torch.cuda.Event(enable_timing=True)


def timed(fn, n=5, skip=2):
    t = 0
    for it in range(n):
        a, b = ev(), ev()
        a.record(st); fn(); b.record(st)
        torch.cuda.synchronize()
        if it >= skip: t += a.elapsed_time(b)
    return t / (n - skip)


def run(wl, tag):
    cap = wl.out_capacity()
    d_tus, d_tuoff, d_par, d_cap = B(wl.tus), B(wl.tu_offsets), B(wl.params), B(cap)
    n_sub = wl.n_sub
    d_co = torch.from_numpy(wl.coeffs).to(dev)
    d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
    d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev); d_bins = torch.zeros_like(d_sizes)
    d_est = torch.zeros(n_sub, dtype=torch.int64, device=dev)
    d_per = torch.zeros(len(wl.tus), dtype=torch.int64, device=dev)
    te = timed(lambda: eng.encode_residual_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_out, d_cap, d_sizes, d_bins))
    bins = int(d_bins.sum().item())
    ts = timed(lambda: eng.estimate_residual_bits_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_est, d_per, d_bins))
    # getCtxInitId inputs + decision: the estimator kernel with per-context counters, then one warp per list
    d_cb = torch.zeros(n_sub * 379, dtype=torch.int32, device=dev)
    d_ctx = torch.zeros(n_sub * 379 * 4, dtype=torch.uint8, device=dev)
    d_choice = torch.zeros(n_sub, dtype=torch.uint8, device=dev)
    tc = timed(lambda: eng.residual_ctx_bins_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_cb, d_bins, ctx_out=d_ctx))
    ti = timed(lambda: eng.ctx_init_id_dev(d_par, d_ctx, d_cb, n_sub, d_choice))
    print(f"{tag}: ctx bin counts {tc:.2f} ms (counted {int(d_cb.sum(dtype=torch.int64).item()) / 1e6:.1f} M context-coded "
          f"bins), init-table choice {ti:.3f} ms, choices {torch.bincount(d_choice.int(), minlength=3).tolist()}")
    trm = int(((wl.params["flags"] & 1) != 0).sum())  # CABAC_SUB_TERMINATE: the coder counts its terminating bin
    assert int(d_bins.sum().item()) + trm == bins
    bits, size = int(d_est.sum().item()) / 32768.0, int(d_sizes.sum().item()) * 8
    print(f"{tag}: {n_sub} lists, {len(wl.tus)} TUs, {bins / 1e6:.1f} Mbins | encode {te:.2f} ms "
          f"({bins / te / 1e6:.2f} Gbins/s) | estimate {ts:.2f} ms ({bins / ts / 1e6:.2f} Gbins/s, "
          f"{len(wl.tus) / ts / 1e3:.2f} MTU/s) | estimated/real bits {bits / size:.4f}")


run(W.frames_as_ctu_substreams(64, 1920, 1080, seed0=1000), "bench batch (CTU substreams)")
run(W.random_substreams(60_000, 77, flags=0, tus_per_sub=(1, 4)), "RDO batch (1-3 TUs per candidate)")
