"""Drop-in proof: the reference's own CABACWriter / CABACReader (compiled unmodified into
oracle/_ref/dropin_test together with tests/dropin/dropin_main.cpp) run on the C++ host shims of
entropy_coding_b200/host/ and must reproduce BinEncoder_Std byte for byte."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "dropin_test")
HOST = os.path.join(ROOT, "entropy_coding_b200", "host")


def test_host_layer_is_reference_independent(tmp_path):
    """cabac_host.hpp (recorder + batches above the C ABI) compiles without any reference header."""
    src = tmp_path / "t.cpp"
    src.write_text('#include "cabac_host.hpp"\nint main() { cabac_b200::StreamRecorder r; r.encodeBinsEP(5, 30);'
                   ' r.encodeRemAbsEP(7, 1, 4, 12); return int(r.cmds.size()) - 4; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I", HOST, "-I", os.path.join(ROOT, "include"),
                           str(src)])


@pytest.mark.gpu
@pytest.mark.parametrize("n_sub,seed", [(96, 7), (400, 11)])
def test_reference_writer_on_gpu_shims(n_sub, seed):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/dropin_test not built (needs the reference tree at build time)")
    r = subprocess.run([BIN, str(n_sub), str(seed)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "DROPIN OK" in r.stdout


@pytest.mark.gpu
def test_slice_reader_shim_on_vtm_slices(tmp_path):
    """GpuSliceReader (host/gpu_slice_reader.hpp, compiled against the unmodified reference headers) decodes the
    VTM-produced fixture slices from their bytes through cabac_decode_slices with host pointers; GpuSliceWriter
    (host/gpu_slice_writer.hpp) codes the records back into the same bytes through cabac_encode_slices."""
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/dropin_test not built (needs the reference tree at build time)")
    import sys
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_syntax as T
    items = T._substreams()
    path = tmp_path / "slices.bin"
    with open(path, "wb") as f:
        f.write(np.uint32(len(items)).tobytes())
        for tag, s, sp in items:
            co = s["coeffs"].astype(np.int64)
            sm = int((np.arange(1, co.size + 1, dtype=np.uint64) * co.astype(np.uint64)).sum(dtype=np.uint64))
            f.write(np.array([sp], T.SLICE_DT).tobytes())
            f.write(np.array([s["qp"]], np.int32).tobytes())
            f.write(np.array([s["init_id"], s["flags"], s["payload"].size, len(s["cus"]), len(s["tus"]), s["num_bins"],
                              co.size], np.uint32).tobytes())
            f.write(np.uint64(sm).tobytes())
            f.write(np.ascontiguousarray(s["ctx_in"].astype(np.uint16)).tobytes())
            f.write(np.ascontiguousarray(s["payload"]).tobytes())
    r = subprocess.run([BIN, "--slices", str(path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DROPIN OK" in r.stdout, r.stdout + r.stderr
    # and GpuSliceWriter (host/gpu_slice_writer.hpp) turns the records back into the payload bytes
    assert "byte-identical with the VTM payloads" in r.stdout, r.stdout
