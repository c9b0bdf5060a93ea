"""Acceptance test of the drop-in boundary: the reference's own demo (examples.cc, UNCHANGED) built
against vacancy_b200/shim/voxel_carver_shim.cc + libvacancy_b200.so must write the same meshes as
the reference's own build of the same demo.  Both binaries are prebuilt under oracle/_ref (they
need the reference tree to compile) and travel to the GPU box; skipped when they did not."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

pytestmark = pytest.mark.gpu


def read_ply(path):
    with open(path) as f:
        lines = f.read().split("\n")
    nv = int([l for l in lines if l.startswith("element vertex")][0].split()[-1])
    nf = int([l for l in lines if l.startswith("element face")][0].split()[-1])
    start = lines.index("end_header") + 1
    verts = np.array([[float(x) for x in l.split()[:3]] for l in lines[start:start + nv]])
    faces = np.array([[int(x) for x in l.split()[1:4]] for l in lines[start + nv:start + nv + nf]])
    return verts.reshape(-1, 3), faces.reshape(-1, 3), lines


def run_demo(binary, workdir, env=None):
    os.makedirs(os.path.join(workdir, "bin"))
    shutil.copytree(os.path.join(REF_DIR, "data"), os.path.join(workdir, "data"))
    for f in os.listdir(os.path.join(workdir, "data")):
        os.chmod(os.path.join(workdir, "data", f), 0o644)
    r = subprocess.run([binary], cwd=os.path.join(workdir, "bin"), capture_output=True, text=True,
                       timeout=600, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return os.path.join(workdir, "data")


def test_examples_cc_unchanged_through_shim(tmp_path):
    ours, theirs = os.path.join(REF_DIR, "examples_b200"), os.path.join(REF_DIR, "examples_ref")
    if not (os.path.exists(ours) and os.path.exists(theirs)):
        pytest.skip("prebuilt demo binaries not present under oracle/_ref")
    a = run_demo(ours, str(tmp_path / "ours"))
    b = run_demo(theirs, str(tmp_path / "ref"))
    diag = 906.0  # |bb_max - bb_min| of examples.cc:87-99
    for i in range(6):
        for name in ("surface_%05d.ply" % i, "surface_nointerp_%05d.ply" % i):
            with open(os.path.join(a, name), "rb") as f1, open(os.path.join(b, name), "rb") as f2:
                assert f1.read() == f2.read(), name  # byte-identical ASCII PLY
        with open(os.path.join(a, "sdf_%05d.png" % i), "rb") as f1, \
                open(os.path.join(b, "sdf_%05d.png" % i), "rb") as f2:
            assert f1.read() == f2.read()
    # voxel meshes: identical faces; vertices within tolerance (the reference translates one cube
    # back and forth, extract_voxel.cc:292-311, so its coordinates drift by < 1e-6)
    va, fa, _ = read_ply(os.path.join(a, "voxel_00005.ply"))
    vb, fb, _ = read_ply(os.path.join(b, "voxel_00005.ply"))
    assert np.array_equal(fa, fb) and va.shape == vb.shape == (28475 * 24, 3)
    assert np.abs(va - vb).max() <= 1e-5 * diag


def test_examples_cc_unchanged_on_several_slabs(tmp_path):
    """The same unchanged demo with VACANCY_B200_DEVICES=0,0,0: the shim then carves the ONE grid as
    three z slabs through vcb_sharded_* (here three virtual slabs on the one GPU of the test box -
    a host with more GPUs lists them instead) and every file must still be byte-identical to what the
    single-carver run writes: meshes, per-view SDF images, and the voxel cubes."""
    ours = os.path.join(REF_DIR, "examples_b200")
    if not os.path.exists(ours):
        pytest.skip("prebuilt demo binary not present under oracle/_ref")
    a = run_demo(ours, str(tmp_path / "slabs"), env={"VACANCY_B200_DEVICES": "0,0,0"})
    b = run_demo(ours, str(tmp_path / "single"))
    names = ["%s_%05d.%s" % (k, i, e) for i in range(6)
             for k, e in (("surface", "ply"), ("surface_nointerp", "ply"), ("sdf", "png"),
                          ("voxel", "ply"))]
    for name in names:
        with open(os.path.join(a, name), "rb") as f1, open(os.path.join(b, name), "rb") as f2:
            assert f1.read() == f2.read(), name


def test_pointer_batch_overload(tmp_path):
    """vacancy_b200::CarveBatch (include/vacancy_b200_batch.h): all six bunny views in one call from
    C++ host code give the reference's final mesh and cube counts."""
    demo = os.path.join(REF_DIR, "batch_demo_b200")
    if not os.path.exists(demo):
        pytest.skip("prebuilt batch demo not present under oracle/_ref")
    r = subprocess.run([demo, os.path.join(REF_DIR, "data") + "/", str(tmp_path)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.split() == ["8672", "17270", "28475"]
    # the PLY written by the reference's own Mesh::WritePly and by the two GPU-side writers
    with open(tmp_path / "surface_reference_writer.ply", "rb") as f:
        want = f.read()
    assert want.startswith(b"ply\nformat ascii 1.0\nelement vertex 8672\n")
    for name in ("surface_gpu_writer.ply", "surface_gpu_direct.ply"):
        with open(tmp_path / name, "rb") as f:
            assert f.read() == want, name
    # the same host program on two z slabs (ExtractIsoSurfaceToPly assembles the slabs' text blocks)
    out2 = tmp_path / "slabs"
    out2.mkdir()
    r = subprocess.run([demo, os.path.join(REF_DIR, "data") + "/", str(out2)], capture_output=True,
                       text=True, timeout=300, env=dict(os.environ, VACANCY_B200_DEVICES="0,0"))
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.split() == ["8672", "17270", "28475"]
    for name in ("surface_reference_writer.ply", "surface_gpu_writer.ply", "surface_gpu_direct.ply"):
        with open(out2 / name, "rb") as f:
            assert f.read() == want, name
