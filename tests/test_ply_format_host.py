"""CPU checks of the PLY text path: the exact "%g" formatter the GPU kernels run
(vacancy_b200/csrc/ply_format.cuh, compiled for the host by tools/ply_format_check.cc) against
snprintf, the oracle's restatement of Mesh::WritePly against the compiled reference, and the
host-only header entry point of the C ABI."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np
import pytest

from oracle import bindings as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_g6_formatter_matches_snprintf(tmp_path):
    exe = str(tmp_path / "ply_format_check")
    subprocess.run(["g++", "-O2", "-fopenmp", "-o", exe, os.path.join(ROOT, "tools", "ply_format_check.cc")],
                   check=True)
    r = subprocess.run([exe, "4000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "mismatches 0" in r.stdout


def _mesh(seed=3, n=4000):
    rng = np.random.default_rng(seed)
    v = (rng.normal(size=(n, 3)) * 10.0 ** rng.uniform(-9, 9, (n, 1))).astype(np.float32)
    v[:6] = [[0, -0.0, 1e-5], [123456.5, 999999.5, 1e6], [1e-4, 9.99999e-5, 100000], [0.5, 2.5, 1000005 / 2],
             [np.inf, -np.inf, np.nan], [1.4e-45, -3.4e38, 8388608]]
    f = rng.integers(0, n, (n // 2, 3)).astype(np.int32)
    f[0] = [0, 2147483647, -1]
    return v, f


def test_oracle_ply_text_equals_the_reference(port):
    if not os.path.exists(B.REF_SO):
        pytest.skip("compiled reference not present")
    ref = B.RefLib()
    v, f = _mesh()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "m.ply")
        assert ref.write_ply(path, v, f) == 0
        with open(path, "rb") as fh:
            want = fh.read()
    assert port.ply_text(v, f) == want


def test_ply_header_entry_point(port):
    from vacancy_b200 import capi
    L = capi.load_library()
    n = C.c_size_t()
    assert L.vcb_ply_header(1234567, 89, None, 0, C.byref(n)) == 0
    buf = C.create_string_buffer(n.value)
    assert L.vcb_ply_header(1234567, 89, buf, n.value, None) == 0
    want = port.ply_text(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32))
    assert buf.raw == want.replace(b"vertex 0", b"vertex 1234567").replace(b"face 0", b"face 89")
    assert L.vcb_ply_header(1, 1, buf, 10, None) != 0  # buffer too small


def test_slab_blocks_assemble_under_one_header(port):
    """sharded.assemble_ply: header for the global counts, every slab's vertex lines, then every
    slab's face lines - checked with the oracle's text cut into two fake slabs (host only)."""
    from vacancy_b200 import sharded
    v, f = _mesh(seed=9, n=300)
    whole = port.ply_text(v, f)
    head_len = len(port.ply_text(v[:0], f[:0]).replace(b"vertex 0", b"vertex 300").replace(
        b"face 0", b"face 150"))
    body = whole[head_len:]
    lines = body.split(b"\n")[:-1]
    vlines, flines = [ln + b"\n" for ln in lines[:300]], [ln + b"\n" for ln in lines[300:]]
    assert len(flines) == 150
    parts = [(b"".join(vlines[:120]), b"".join(flines[:40])),
             (b"".join(vlines[120:]), b"".join(flines[40:]))]
    assert sharded.assemble_ply((300, 150), parts) == whole
