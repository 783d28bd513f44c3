"""include/b200ivf.hpp (the header-only C++ face of the C-ABI) compiled with g++ and
driven the way the fork's benchmark drives faiss::IndexIVFFlat. CPU: it compiles, links against
the library and reports the missing device through the exception path. GPU: its results equal the
ctypes mirror's on the same inputs."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "hnsw_ivf_b200", "lib")


def _build(tmp_path):
    exe = str(tmp_path / "hpp_drive")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "hpp_drive.cpp"), "-L", LIBDIR,
                           "-lb200ivf", "-Wl,-rpath," + LIBDIR, "-o", exe])
    return exe


def _inputs(tmp_path, metric):
    from helpers import clustered, sample_centroids
    d, nb, nlist, nq, k, nprobe = 32, 6000, 32, 40, 10, 6
    xb = clustered(nb, d, 4 * nlist, 1, normalize=metric == 0)
    xq = clustered(nq, d, 4 * nlist, 2, normalize=metric == 0)
    cent = sample_centroids(xb, nlist)
    return (d, nb, nlist, nq, k, nprobe), cent, xb, xq


def _write(path, shape, metric, radius, cent, xb, xq):
    with open(path, "wb") as f:
        f.write(np.array(list(shape) + [metric], np.int64).tobytes())
        f.write(np.float32(radius).tobytes())
        for a in (cent, xb, xq):
            f.write(np.ascontiguousarray(a, np.float32).tobytes())


def test_header_compiles_and_reports_no_device(tmp_path):
    import torch
    exe = _build(tmp_path)
    shape, cent, xb, xq = _inputs(tmp_path, 1)
    _write(tmp_path / "in.bin", shape, 1, 1.0, cent, xb, xq)
    r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    if not torch.cuda.is_available():
        assert "NO_DEVICE code=-2" in r.stdout  # FaissException's code, c_api/error_c.h:26
    else:
        assert r.stdout.startswith("OK")


@pytest.mark.gpu
@pytest.mark.parametrize("metric", [1, 0])
def test_header_results_equal_ctypes(cuda_device, tmp_path, metric):
    import hnsw_ivf_b200 as b2
    exe = _build(tmp_path)
    shape, cent, xb, xq = _inputs(tmp_path, metric)
    d, nb, nlist, nq, k, nprobe = shape
    g = b2.IndexIVFFlat(d, nlist, metric)
    g.set_centroids(cent)
    g.add(xb)
    g.nprobe = nprobe
    D, I = g.search(xq, k)
    radius = float(np.median(D[:, k // 2]))
    _write(tmp_path / "in.bin", shape, metric, radius, cent, xb, xq)
    r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "OK ntotal=%d trained=1" % nb in r.stdout, r.stdout + r.stderr
    assert "THROWS code=-2" in r.stdout and "k > 0" in r.stdout
    assert "SHARDED_OK" in r.stdout, r.stdout  # b200::IndexShardsIVFFlat, world of one rank
    buf = open(tmp_path / "out.bin", "rb").read()
    off = 0

    def take(dtype, count):
        nonlocal off
        a = np.frombuffer(buf, dtype=dtype, count=count, offset=off)
        off += a.nbytes
        return a
    Dc, Ic = take(np.float32, nq * k).reshape(nq, k), take(np.int64, nq * k).reshape(nq, k)
    Dsc, Isc = take(np.float32, nq * k).reshape(nq, k), take(np.int64, nq * k).reshape(nq, k)
    lims = take(np.uint64, nq + 1)
    rI = take(np.int64, int(lims[-1]))
    rD = take(np.float32, int(lims[-1]))
    assert off == len(buf)
    assert np.array_equal(I, Ic) and np.array_equal(D, Dc)
    p = b2.SearchParametersIVF(sel=b2.IDSelectorRange(0, nb // 2), nprobe=nprobe)
    Ds, Is = g.search(xq, k, params=p)
    assert np.array_equal(Is, Isc) and np.array_equal(Ds, Dsc) and Isc.max() < nb // 2
    l2, D2, I2 = g.range_search(xq, radius)
    assert np.array_equal(l2, lims) and np.array_equal(I2, rI) and np.array_equal(D2, rD)
