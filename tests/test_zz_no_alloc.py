"""cabac_decode_residual_static_dev: the allocation-free residual decode (context rows in a static device array;
no counterpart in the reference). Collected last on purpose:
a failure here must not hide the other tests.

Checked through the C ABI against cabac_decode_residual_dev on the same device buffers, bit-exact (the tolerance
every integer path in this suite uses), on two calls in a row that share the static array."""
import ctypes as C

import numpy as np
import pytest

from entropy_coding_b200 import workload as W
from entropy_coding_b200.types import CTX_DT, ERR_ARG, NUM_CTX, STATIC_MAX_SUBSTREAMS, SUB_CTX_IN, SUB_DEP_QUANT
from helpers import T

pytestmark = pytest.mark.gpu


def _device_batch(torch, eng, wl):
    """Encodes wl with the ordinary entry points and leaves everything a decode call needs on the device."""
    dev = torch.device("cuda", 0)
    r = eng.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, wl.out_capacity(bytes_per_coeff=5.0))
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)
    n, n_tu = wl.n_sub, len(wl.tus)
    b = {"n": n, "enc": r,
         "in": up(r["packed"]), "ent": up(r["entry_points"].astype(np.uint64)), "tus": up(wl.tus),
         "tu_off": up(wl.tu_offsets.astype(np.uint64)), "params": up(wl.params)}
    for tag, fill in (("a", 0x5A), ("b", 0xC3)):  # outputs of the existing entry point / of the new one: different junk
        b[tag] = {"coeffs": torch.full((wl.coeffs.size,), fill * 0x010101, dtype=torch.int32, device=dev),
                  "num_bins": torch.full((n,), fill, dtype=torch.int32, device=dev),
                  "bits_read": torch.full((n,), fill, dtype=torch.int32, device=dev),
                  "tu_flags": torch.full((n_tu,), fill, dtype=torch.uint8, device=dev),
                  "ctx_out": torch.full((n * NUM_CTX * CTX_DT.itemsize,), fill, dtype=torch.uint8, device=dev),
                  "status": torch.full((n,), fill, dtype=torch.int32, device=dev)}
    return b


def _run_pair(torch, eng, wl):
    b = _device_batch(torch, eng, wl)
    torch.cuda.synchronize()
    o = b["a"]
    eng.set_decoder_mode(1)  # the existing entry point on its default (lockstep) kernel
    eng.decode_residual_dev(b["in"], b["ent"], b["tus"], b["tu_off"], b["params"], b["n"], o["coeffs"], o["num_bins"],
                            o["bits_read"], o["tu_flags"], None, o["ctx_out"], o["status"])
    eng.sync()
    free0 = torch.cuda.mem_get_info()[0]
    launches0 = eng.lib.cabac_launch_count(eng.h)
    o = b["b"]
    eng.decode_residual_static_dev(b["in"], b["ent"], b["tus"], b["tu_off"], b["params"], b["n"], o["coeffs"],
                                   o["num_bins"], o["bits_read"], o["tu_flags"], None, o["ctx_out"], o["status"])
    eng.sync()
    assert torch.cuda.mem_get_info()[0] == free0, "the allocation-free call changed the device's free memory"
    assert eng.lib.cabac_launch_count(eng.h) - launches0 == 2
    for k in ("coeffs", "num_bins", "bits_read", "tu_flags", "ctx_out", "status"):
        assert torch.equal(b["a"][k], b["b"][k]), k
    assert int(b["b"]["status"].abs().max()) == 0
    assert (b["b"]["coeffs"].cpu().numpy() == wl.coeffs).all()
    assert (b["b"]["num_bins"].cpu().numpy().astype(np.uint32) == b["enc"]["num_bins"]).all()
    return b


def test_static_decode_matches_existing_entry_point_twice():
    torch = pytest.importorskip("torch")
    import entropy_coding_b200 as E
    eng = E.CabacEngine(0)
    try:
        # first call: 3 000 substreams; second call: fewer, different substreams over the same (now dirty) rows
        wl1 = W.random_substreams(3000, seed=7001, tus_per_sub=(1, 10), ts_fraction=0.25, bdpcm_fraction=0.05,
                                  sbt_fraction=0.05)
        wl2 = W.random_substreams(700, seed=7002, flags=T | SUB_DEP_QUANT, tus_per_sub=(2, 24), ts_fraction=0.1)
        _run_pair(torch, eng, wl1)
        _run_pair(torch, eng, wl2)
        # and the first batch again after the second: same answer as before
        _run_pair(torch, eng, wl1)
    finally:
        eng.close()


def test_static_decode_context_handover_and_empty_batch():
    """Handed-in context models (CABAC_SUB_CTX_IN) go through the static rows like the initialised ones."""
    torch = pytest.importorskip("torch")
    import entropy_coding_b200 as E
    eng = E.CabacEngine(0)
    try:
        dev = torch.device("cuda", 0)
        wl = W.random_substreams(400, seed=7003, tus_per_sub=(1, 8), ts_fraction=0.2)
        first = eng.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs,
                                           wl.out_capacity(bytes_per_coeff=5.0), want_ctx=True)
        params = wl.params.copy()
        params["flags"] |= SUB_CTX_IN
        r = eng.encode_residual_packed(wl.tus, wl.tu_offsets, params, wl.coeffs, wl.out_capacity(bytes_per_coeff=5.0),
                                       ctx_in=first["ctx_out"])
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).to(dev)
        d_in, d_ent, d_tus = up(r["packed"]), up(r["entry_points"].astype(np.uint64)), up(wl.tus)
        d_off, d_par, d_ctx = up(wl.tu_offsets.astype(np.uint64)), up(params), up(first["ctx_out"])
        outs = []
        for call in (eng.decode_residual_dev, eng.decode_residual_static_dev, eng.decode_residual_static_dev):
            co = torch.full((wl.coeffs.size,), 77, dtype=torch.int32, device=dev)
            nb = torch.zeros(wl.n_sub, dtype=torch.int32, device=dev)
            cx = torch.zeros(wl.n_sub * NUM_CTX * CTX_DT.itemsize, dtype=torch.uint8, device=dev)
            torch.cuda.synchronize()
            call(d_in, d_ent, d_tus, d_off, d_par, wl.n_sub, co, nb, None, None, d_ctx, cx, None)
            eng.sync()
            outs.append((co, nb, cx))
        for o in outs[1:]:
            assert all(torch.equal(x, y) for x, y in zip(outs[0], o))
        assert (outs[1][0].cpu().numpy() == wl.coeffs).all()
        # an empty batch is a no-op
        eng.decode_residual_static_dev(None, None, None, None, None, 0, None)
    finally:
        eng.close()


def test_static_decode_refuses_batches_beyond_the_maximum():
    pytest.importorskip("torch")
    import entropy_coding_b200 as E
    eng = E.CabacEngine(0)
    try:
        one = C.c_void_p(256)  # never dereferenced: the size check comes first
        rc = eng.lib.cabac_decode_residual_static_dev(eng.h, one, one, one, one, one, STATIC_MAX_SUBSTREAMS + 1, None, one,
                                                      None, None, None, None, None)
        assert rc == ERR_ARG
        msg = eng.lib.cabac_last_error(eng.h).decode()
        assert "CABAC_STATIC_MAX_SUBSTREAMS" in msg and str(STATIC_MAX_SUBSTREAMS) in msg
        with pytest.raises(E.CabacError):
            eng.decode_residual_static_dev(256, 256, 256, 256, 256, STATIC_MAX_SUBSTREAMS + 1, 256)
    finally:
        eng.close()


def test_static_decode_at_the_maximum():
    """Exactly CABAC_STATIC_MAX_SUBSTREAMS substreams: the last row of the array is used."""
    torch = pytest.importorskip("torch")
    import entropy_coding_b200 as E
    eng = E.CabacEngine(0)
    try:
        wl = W.random_substreams(STATIC_MAX_SUBSTREAMS, seed=7004, tus_per_sub=(1, 3))
        _run_pair(torch, eng, wl)
    finally:
        eng.close()
