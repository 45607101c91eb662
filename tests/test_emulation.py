"""Start-code emulation counting / prevention: oracle vs the reference's own
OutputBitstream::countStartCodeEmulations on CPU, the CUDA kernels vs the oracle on the GPU."""
import numpy as np
import pytest


def cases():
    rng = np.random.default_rng(17)
    out = [np.zeros(0, np.uint8), np.array([0], np.uint8), np.array([0, 0], np.uint8), np.array([0, 0, 0], np.uint8),
           np.array([0, 0, 3], np.uint8), np.array([0, 0, 4], np.uint8), np.array([1, 0, 0, 1, 0, 0, 0, 0, 2, 0, 0], np.uint8),
           np.zeros(37, np.uint8), np.zeros(4100, np.uint8)]
    for n in (5, 64, 1000, 4096, 4097, 20000):
        out.append(rng.integers(0, 256, n).astype(np.uint8))                       # arithmetic-coder-like
        out.append(rng.choice(np.array([0, 0, 0, 1, 2, 3, 4, 255], np.uint8), n))  # dense in emulations
    return out


def test_oracle_count_matches_reference(oracle, ref):
    for d in cases():
        assert oracle.count_emulations(d) == ref.count_emulations(d), d[:16]


def test_oracle_insert_is_consistent_with_count(oracle):
    for d in cases():
        out = oracle.insert_emulation_prevention(d)
        trailing = 1 if d.size and d[-1] == 0 else 0
        assert out.size == d.size + oracle.count_emulations(d) + trailing
        # the NAL payload never contains 00 00 0x (x <= 2) and removing the 03s restores the RBSP
        b = out.tobytes()
        assert b"\x00\x00\x00" not in b and b"\x00\x00\x01" not in b and b"\x00\x00\x02" not in b
        back, z = bytearray(), 0
        for v in out[:out.size - trailing]:
            if z == 2 and v == 3:
                z = 0
                continue
            z = z + 1 if v == 0 else 0
            back.append(v)
        assert bytes(back) == d.tobytes()


@pytest.mark.gpu
def test_gpu_emulation_kernels(engine, oracle):
    cs = [c for c in cases() if c.size]
    data = np.concatenate(cs)
    offsets = np.concatenate([[0], np.cumsum([c.size for c in cs])]).astype(np.uint64)
    counts = engine.count_emulations(data, offsets)
    assert counts.tolist() == [oracle.count_emulations(c) for c in cs]
    for c in cs + [data]:
        got = engine.insert_emulation_prevention(c)
        assert got.tobytes() == oracle.insert_emulation_prevention(c).tobytes()


@pytest.mark.gpu
def test_gpu_entry_point_offsets_with_emulation(engine, oracle):
    """Entry-point offsets of real substreams as VTM signals them (EncGOP.cpp:3811-3830):
    substream size + its emulation count; the NAL payload grows by exactly the sum."""
    from entropy_coding_b200 import workload as W
    wl = W.random_substreams(300, 5, tus_per_sub=(3, 10))
    r = engine.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, wl.out_capacity(bytes_per_coeff=5.0))
    counts = engine.count_emulations(r["packed"], r["entry_points"])
    nal = engine.insert_emulation_prevention(r["packed"])
    assert nal.size == r["packed"].size + int(counts.sum())
    assert nal.tobytes() == oracle.insert_emulation_prevention(r["packed"]).tobytes()
    for i in range(0, 300, 37):
        a, b = int(r["entry_points"][i]), int(r["entry_points"][i + 1])
        assert int(counts[i]) == oracle.count_emulations(r["packed"][a:b])
