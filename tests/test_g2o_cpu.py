"""g2o / TORO text format: the product's parser and writer, exercised on the CPU through the
test-only hook gs_host_g2o (same C++ as gs_load_g2o / gs_save_g2o)."""
import ctypes as C

import numpy as np
import pytest

from graphslam_b200 import _lib, synth
from graphslam_b200._lib import c_double_p, c_u64_p


@pytest.fixture(scope="module")
def lib():
    return _lib.load_debug()


def _parse(lib, path, out_path=None):
    cnt = (C.c_size_t * 5)()
    rc = lib.gs_host_g2o(str(path).encode(), None, cnt, None, None, None, None, None, None, None, None, None)
    if rc != 0:
        return rc, None
    nv, ne, npr = cnt[0], cnt[1], cnt[2]
    a = dict(vkey=np.zeros(nv, np.uint64), vxyt=np.zeros((nv, 3)), e1=np.zeros(ne, np.uint64), e2=np.zeros(ne, np.uint64),
             ez=np.zeros((ne, 3)), einfo=np.zeros((ne, 6)), pkey=np.zeros(npr, np.uint64), pxyt=np.zeros((npr, 3)),
             pinfo=np.zeros((npr, 6)))
    u = lambda x: x.ctypes.data_as(c_u64_p)
    d = lambda x: x.ctypes.data_as(c_double_p)
    rc = lib.gs_host_g2o(str(path).encode(), str(out_path).encode() if out_path else None, cnt, u(a["vkey"]), d(a["vxyt"]),
                         u(a["e1"]), u(a["e2"]), d(a["ez"]), d(a["einfo"]), u(a["pkey"]), d(a["pxyt"]), d(a["pinfo"]))
    a["counts"] = [int(c) for c in cnt]
    return rc, a


def test_reads_g2o_and_toro_tags(lib, tmp_path):
    f = tmp_path / "mixed.g2o"
    f.write_text(
        "# a comment line\n"
        "VERTEX_SE2 0 0 0 0\n"
        "VERTEX_SE2 7 1.5 -2.25 0.1   # trailing comment\n"
        "VERTEX2 12 3e-1 4 -3.0\n"
        "\n"
        "EDGE_SE2 0 7 1 2 0.3 11 12 13 22 23 33\n"
        "EDGE2 7 12 -1 0.5 0.25 11 12 22 33 13 23\n"
        "EDGE_PRIOR_SE2 0 0 0 0 1000 0 0 1000 0 1000\n"
        "FIX 0\n"
        "VERTEX_XY 99 1 2\n")
    rc, a = _parse(lib, f)
    assert rc == 0
    assert a["counts"] == [3, 2, 1, 1, 1]
    assert a["vkey"].tolist() == [0, 7, 12]
    assert np.array_equal(a["vxyt"], [[0, 0, 0], [1.5, -2.25, 0.1], [0.3, 4, -3.0]])
    assert a["e1"].tolist() == [0, 7] and a["e2"].tolist() == [7, 12]
    assert np.array_equal(a["ez"], [[1, 2, 0.3], [-1, 0.5, 0.25]])
    # both tags end up as I11 I12 I13 I22 I23 I33
    assert np.array_equal(a["einfo"], [[11, 12, 13, 22, 23, 33], [11, 12, 13, 22, 23, 33]])
    assert a["pkey"].tolist() == [0] and np.array_equal(a["pinfo"], [[1000, 0, 0, 1000, 0, 1000]])


def test_write_then_read_is_bit_exact(lib, tmp_path):
    g = synth.config("C1", n=400)
    info = np.array([np.linalg.inv(q.reshape(3, 3))[np.triu_indices(3)] for q in g.cov])
    src = tmp_path / "a.g2o"
    with open(src, "w") as fh:
        for k, p in zip(g.keys, g.init):
            fh.write("VERTEX_SE2 %d %s %s %s\n" % (k, repr(float(p[0])), repr(float(p[1])), repr(float(p[2]))))
        for a_, b_, z, o in zip(g.k1, g.k2, g.z, info):
            fh.write("EDGE_SE2 %d %d " % (a_, b_) + " ".join(repr(float(v)) for v in list(z) + list(o)) + "\n")
    out = tmp_path / "b.g2o"
    rc, a = _parse(lib, src, out)
    assert rc == 0 and a["counts"][:3] == [len(g.keys), len(g.k1), 0]
    assert np.array_equal(a["vxyt"], g.init) and np.array_equal(a["ez"], g.z) and np.array_equal(a["einfo"], info)
    rc, b = _parse(lib, out)
    assert rc == 0
    for k in ("vkey", "vxyt", "e1", "e2", "ez", "einfo"):
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("text", ["VERTEX_SE2 1 2 3\n", "EDGE_SE2 0 1 1 2 3 1 0 0 1 0\n", "VERTEX_SE2 x 1 2 3\n",
                                  "EDGE_SE2 0 -1 1 2 3 1 0 0 1 0 1\n"])
def test_malformed_lines_are_errors(lib, tmp_path, text):
    f = tmp_path / "bad.g2o"
    f.write_text(text)
    rc, _ = _parse(lib, f)
    assert rc == _lib.GS_EINVAL


def test_missing_file_is_an_error(lib, tmp_path):
    rc, _ = _parse(lib, tmp_path / "nope.g2o")
    assert rc == _lib.GS_EINVAL
