"""Substream assembly on the decode side and the pin of the emulation
prevention oracle to VTM's own NAL code.

  * oracle insert / remove == VTM NALwrite.cpp / NALread.cpp compiled unmodified (oracle/_ref/libnal_ref.so)
  * the real str.bin of the VTM fixtures: EPB removal of the slice NAL units exposes the recorded payloads
    verbatim; for the WPP streams the slice header's tail bits equal sh_entry_offset_len_minus1 +
    sh_entry_point_offset_minus1[] computed from the substream sizes and emulation counts
    (VLCWriter.cpp:2833-2866), and the decoder-side range recovery (VLCReader.cpp:4346-4381) finds every row
  * -m gpu: the device kernels against the oracle
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from vtm_streams import load_fixture  # noqa: E402

FIX = [n for n in ("vtm_qp51.npz", "vtm_qp32.npz", "vtm_wpp_qp32.npz") if os.path.exists(os.path.join(HERE, "golden", n))]


@pytest.fixture(scope="module")
def nalref():
    from oracle.checker import NalRef
    if not NalRef.available():
        pytest.skip("oracle/_ref/libnal_ref.so not built (VTM tree absent)")
    return NalRef()


def zero_heavy(rng, n):
    return rng.choice(np.array([0, 0, 0, 1, 2, 3, 4, 255], np.uint8), size=n, p=[.2, .2, .2, .1, .08, .08, .07, .07])


def test_oracle_emulation_prevention_pinned_to_vtm(oracle, nalref):
    rng = np.random.default_rng(1)
    n_epb = 0
    for it in range(2000):
        r = zero_heavy(rng, int(rng.integers(1, 300)))
        a = oracle.insert_emulation_prevention(r)
        assert a.tobytes() == nalref.insert(r).tobytes()  # writeNaluContent, NALwrite.cpp:67-126
        vcl = it % 2
        nal = np.concatenate([np.array([0, (1 << 3) if vcl else (14 << 3)], np.uint8), a])
        ro, rv = oracle.remove_emulation_prevention(nal, vcl=vcl), nalref.remove(nal)  # NALread.cpp:59-106
        assert ro is not None and rv is not None
        assert ro[0].tobytes() == rv[0].tobytes() and (ro[1] == rv[1]).all()
        n_epb += len(ro[1])
        if r[-1] != 0:
            assert ro[0][2:].tobytes() == r.tobytes()
    assert n_epb > 5000


def test_oracle_refuses_malformed_payloads(oracle):
    """The CHECKs of convertPayloadToRBSP: 00 00 0x with x < 3, a byte > 3 after an EPB, a trailing zero."""
    for bad in ([9, 9, 0, 0, 1, 5], [9, 9, 0, 0, 0, 5], [9, 9, 0, 0, 3, 7, 1], [9, 9, 5, 0]):
        assert oracle.remove_emulation_prevention(np.asarray(bad, np.uint8)) is None
    ok = oracle.remove_emulation_prevention(np.asarray([9, 9, 0, 0, 3, 1, 0, 0, 3], np.uint8))
    assert ok[0].tolist() == [9, 9, 0, 0, 1, 0, 0] and ok[1].tolist() == [4, 8]


def split_nal_units(blob: bytes):
    """Annex-B byte stream -> NAL units (start code prefixes 00 00 01 / 00 00 00 01 removed)."""
    out, i, n = [], 0, len(blob)
    starts = []
    while i + 2 < n:
        if blob[i] == 0 and blob[i + 1] == 0 and blob[i + 2] == 1:
            starts.append(i + 3)
            i += 3
        else:
            i += 1
    for k, s in enumerate(starts):
        e = (starts[k + 1] - 3) if k + 1 < len(starts) else n
        while e > s and blob[e - 1] == 0:  # trailing_zero_8bits / the zero_byte of the next start code
            e -= 1
        out.append(np.frombuffer(blob[s:e], np.uint8))
    return out


def bits_of(data: np.ndarray) -> str:
    return "".join(f"{b:08b}" for b in data.tolist())


def slices_of(name, oracle):
    """(rbsp, epb, [substreams of this slice]) per coded slice NAL of the fixture's str.bin."""
    subs, extra = load_fixture(name)
    nals = [u for u in split_nal_units(extra["str_bin"].tobytes()) if (u[1] >> 3) <= 11]  # VCL: slices
    per = len(subs) // len(nals)
    assert per * len(nals) == len(subs)
    res = []
    for k, u in enumerate(nals):
        rbsp, epb = oracle.remove_emulation_prevention(u, vcl=True)
        res.append((u, rbsp, epb, subs[k * per:(k + 1) * per]))
    return res


@pytest.mark.parametrize("name", FIX)
def test_str_bin_slices_decode_side(oracle, name):
    for nal, rbsp, epb, rows in slices_of(name, oracle):
        data = np.concatenate([s["payload"] for s in rows])
        blob = rbsp.tobytes()
        at = blob.rfind(data.tobytes())
        assert at >= 0 and at + data.size == rbsp.size, "the slice data is the tail of the slice NAL's RBSP"
        # encoder side: what EncSlice / HLSWriter signal for these substreams
        sizes = np.asarray([s["payload"].size for s in rows], np.uint32)
        emu = np.asarray([oracle.count_emulations(s["payload"]) for s in rows], np.uint32)
        ln, off = oracle.entry_point_syntax(sizes, emu)
        if len(rows) > 1:
            # slice header tail: ue(v) sh_entry_offset_len_minus1, u(len) offsets, then byte_alignment (1 0*)
            ue = "0" * (len(bin(ln + 1)) - 3) + bin(ln + 1)[2:]
            want = ue + "".join(format(int(o), f"0{ln + 1}b") for o in off) + "1"
            hdr = bits_of(rbsp[:at])
            tail = hdr.rstrip("0")
            assert tail.endswith(want), f"{name}: slice header does not end with the computed entry points"
            # decoder side: ranges recovered from the signalled offsets + EPB locations
            rng_ = oracle.substream_ranges(off, epb, at, rbsp.size)
            pos = at
            for i, s in enumerate(rows):
                assert int(rng_[i]) == pos and int(rng_[i + 1]) == pos + s["payload"].size
                pos += s["payload"].size
        # insertion over the whole RBSP gives back the NAL unit VTM wrote
        assert oracle.insert_emulation_prevention(rbsp).tobytes() == nal.tobytes()


# ------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_gpu_remove_emulation_prevention(engine, oracle):
    rng = np.random.default_rng(7)
    for it, n in enumerate([1, 2, 3, 7, 100, 4095, 4096, 4097, 70000, 1 << 20]):
        r = zero_heavy(rng, n) if it % 3 else rng.integers(0, 256, n).astype(np.uint8)
        nal = np.concatenate([np.array([0, 8], np.uint8), oracle.insert_emulation_prevention(r)])
        for vcl in (False, True):
            want = oracle.remove_emulation_prevention(nal, vcl=vcl)
            got = engine.remove_emulation_prevention(nal, vcl=vcl)
            assert got["rbsp"].tobytes() == want[0].tobytes() and (got["epb_locations"] == want[1]).all()
            assert got["violations"] == 0
    # a round trip through the device insertion kernel as well
    r = zero_heavy(rng, 300000)
    r[-1] = 9
    back = engine.remove_emulation_prevention(engine.insert_emulation_prevention(r))
    assert back["rbsp"].tobytes() == r.tobytes()
    for bad in ([9, 9, 0, 0, 1, 5], [9, 9, 0, 0, 0, 5], [9, 9, 0, 0, 3, 7, 1], [9, 9, 5, 0]):
        assert engine.remove_emulation_prevention(np.asarray(bad, np.uint8), raise_on_violation=False)["violations"] > 0


@pytest.mark.gpu
@pytest.mark.parametrize("name", FIX)
def test_gpu_slice_disassembly_of_str_bin(engine, oracle, name):
    """NAL unit of a real slice -> RBSP -> substream ranges on the device; the ranges feed cabac_decode_*."""
    for nal, rbsp, epb, rows in slices_of(name, oracle):
        got = engine.remove_emulation_prevention(nal, vcl=True)
        assert got["rbsp"].tobytes() == rbsp.tobytes() and (got["epb_locations"] == epb).all()
        sizes = np.asarray([s["payload"].size for s in rows], np.uint32)
        emu = engine.count_emulations(np.concatenate([s["payload"] for s in rows]),
                                      np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64))
        ln, off = engine.entry_point_syntax(sizes, emu)
        assert (ln, off.tolist()) == (oracle.entry_point_syntax(sizes, emu)[0], oracle.entry_point_syntax(sizes, emu)[1].tolist())
        at = rbsp.size - int(sizes.sum())
        r = engine.substream_ranges(off, epb, at, rbsp.size)
        assert (r == oracle.substream_ranges(off, epb, at, rbsp.size)).all()
        for i, s in enumerate(rows):
            assert rbsp[int(r[i]):int(r[i + 1])].tobytes() == s["payload"].tobytes()
