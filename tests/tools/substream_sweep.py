"""Dev tool: BASELINE configs[4] - the substream-count sweep (10^4, 10^5, 10^6 substreams per launch).

A seeded base batch of 10^4 short substreams (2-6 TUs, mixed QP presets) is tiled on the device to
10^5 and 10^6 substreams (every copy owns its own coefficient tile, output slab and decode target), then
encode (AUTO: two-stage from 32 768 substreams) + pack + decode are timed with CUDA events and the round
trip is checked on every tile.  usage: substream_sweep.py [max_tiles=100]"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

max_tiles = int(sys.argv[1]) if len(sys.argv) > 1 else 100
base = W.random_substreams(10_000, seed=42 + 10_000, tus_per_sub=(2, 6))
dev = torch.device("cuda", 0)
eng = E.CabacEngine(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); eng.set_stream(st.cuda_stream)
B = lambda a: torch.from_numpy(np.frombuffer(np.ascontiguousarray(a).tobytes(), np.uint8).copy()).to(dev)
cap1 = base.out_capacity()
nco1, ntu1 = base.coeffs.size, len(base.tus)
for tiles in (1, 10, 100):
    if tiles > max_tiles: break
    n_sub = base.n_sub * tiles
    tus = np.tile(base.tus, tiles)
    tus["coeff_offset"] += np.repeat(np.arange(tiles, dtype=np.uint64) * np.uint64(nco1), ntu1)
    tu_off = np.concatenate([[0], (np.arange(tiles)[:, None] * ntu1 + base.tu_offsets[1:].astype(np.int64)[None, :]).ravel()]).astype(np.uint64)
    params = np.tile(base.params, tiles)
    cap = np.concatenate([[0], (np.arange(tiles)[:, None] * int(cap1[-1]) + cap1[1:].astype(np.int64)[None, :]).ravel()]).astype(np.uint64)
    d_tus, d_tuoff, d_par, d_cap = B(tus), B(tu_off), B(params), B(cap)
    d_base = torch.from_numpy(base.coeffs).to(dev)
    d_co = d_base.repeat(tiles)
    d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
    d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev); d_bins = torch.zeros_like(d_sizes)
    d_packed = torch.empty_like(d_out); d_entry = torch.zeros(n_sub + 1, dtype=torch.int64, device=dev)
    d_dec = torch.empty_like(d_co)
    te, td = [], []
    for it in range(4):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record(st)
        eng.encode_residual_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_out, d_cap, d_sizes, d_bins)
        eng.pack_dev(d_out, d_cap, d_sizes, n_sub, d_packed, d_entry)
        ev[1].record(st)
        d_dec.fill_(0x5a5a5a5a)
        ev[2].record(st)
        eng.decode_residual_dev(d_packed, d_entry, d_tus, d_tuoff, d_par, n_sub, d_dec, d_bins)
        ev[3].record(st)
        torch.cuda.synchronize()
        if it: te.append(ev[0].elapsed_time(ev[1])); td.append(ev[2].elapsed_time(ev[3]))
    ok = bool((d_dec.view(tiles, nco1) == d_base[None, :]).all().item())
    bins = int(d_bins.sum(dtype=torch.int64).item()); nbytes = int(d_sizes.sum(dtype=torch.int64).item())
    same = bool((d_sizes.view(tiles, -1) == d_sizes[:base.n_sub][None, :]).all().item())
    print(json.dumps({"substreams": n_sub, "tus": int(len(tus)), "coeff_positions": int(d_co.numel()), "bins": bins,
                      "bitstream_bytes": nbytes, "encode_pack_ms": round(min(te), 3), "decode_ms": round(min(td), 3),
                      "encode_gbins_s": round(bins / min(te) / 1e6, 2), "decode_gbins_s": round(bins / min(td) / 1e6, 2),
                      "round_trip_ok": ok, "tiles_identical": same}), flush=True)
    del d_tus, d_tuoff, d_par, d_cap, d_co, d_out, d_sizes, d_bins, d_packed, d_entry, d_dec
    torch.cuda.empty_cache()
