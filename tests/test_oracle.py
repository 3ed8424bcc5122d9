"""The oracle (oracle/estep_oracle.c) against the golden vectors -- outputs of the reference's
own weaklabel.cpp (tests/golden/make_golden.py) -- and, where oracle/_ref is built, against the
reference binary live.  This is what "parity PINNED" in the oracle's header rests on."""
import os
import subprocess
import sys

import numpy as np
import pytest

import cases
from oracle.oracle import Oracle, Reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALL = [c["name"] for c in cases.CASES]
# the CPU suite must finish in minutes: the three biggest cases (30x321^2, 4x513^2) stay, they
# cost a few seconds each through the C restatement
CPU_CASES = ALL


@pytest.mark.parametrize("name", CPU_CASES)
def test_oracle_matches_golden(name, golden):
    meta, arrays = golden
    f, l, p = cases.build_case(name)
    out = Oracle(1).e_step(f, l, **p)
    lab = np.argmax(out, axis=3).astype(np.uint8)
    assert cases.sha1(out) == meta[name]["map_sha1"]
    assert cases.sha1(lab) == meta[name]["labels_sha1"]
    if name + "/labels" in arrays.files:
        assert np.array_equal(lab, arrays[name + "/labels"])
    if name + "/map" in arrays.files:
        assert out.tobytes() == arrays[name + "/map"].tobytes()


def test_oracle_argmax_is_first_max(oracle):
    x = np.zeros((1, 2, 2, 21), np.float32)
    x[0, 0, 1, 5] = 1; x[0, 0, 1, 9] = 1
    x[0, 1, 0, 20] = 2
    lab = oracle.argmax_labels(x)
    assert lab.tolist() == [[[0, 5], [20, 0]]]
    assert np.array_equal(lab, np.argmax(x, axis=3))


def test_oracle_rng_is_glibc_rand():
    """oracle_rand == libc rand() after srand(seed), for a few seeds (fresh subprocess: the
    global stream must not be disturbed by other tests)."""
    code = (
        "import ctypes,sys;sys.path.insert(0,%r);from oracle.oracle import Oracle;"
        "libc=ctypes.CDLL('libc.so.6');ok=True\n"
        "for seed in (1,2,12345,0,2**31-1):\n"
        "    o=Oracle(seed);libc.srand(seed)\n"
        "    ok&=all(o.rand()==libc.rand() for _ in range(2000))\n"
        "o=Oracle(1)\n"
        "print('OK' if ok else 'BAD')" % ROOT)
    out = subprocess.check_output([sys.executable, "-c", code], text=True)
    assert out.strip().endswith("OK")


def test_oracle_first_draws_unseeded_process():
    """A process that never called srand sees seed 1: first rand() values are well known."""
    o = Oracle(1)
    assert [o.rand() for _ in range(3)] == [1804289383, 846930886, 1681692777]


@pytest.mark.skipif(not Reference.available("ref"), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", ["p5_seq3", "ties_int", "nosuppress_p3", "mag_big", "stride_padpix",
                                  "c2_41_b30"])
def test_oracle_matches_reference_live(name):
    """Run the reference's own code in a fresh subprocess (known rand() origin)."""
    code = (
        "import sys;sys.path.insert(0,%r);sys.path.insert(0,%r)\n"
        "import numpy as np,cases\n"
        "from oracle.oracle import Reference,Oracle\n"
        "f,l,p=cases.build_case(%r)\n"
        "want=Oracle(1).e_step(f,l,**p)\n"
        "ref=Reference('ref')\n"
        "for b in range(f.shape[0]): ref.e_step_image_inplace(f[b:b+1],l[b:b+1],**p)\n"
        "print('OK' if np.ascontiguousarray(f).tobytes()==want.tobytes() else 'BAD')"
        % (ROOT, os.path.join(ROOT, "tests"), name))
    out = subprocess.check_output([sys.executable, "-c", code], text=True)
    assert out.strip().endswith("OK")


def test_oracle_rejects_empty_label_set(oracle):
    f = np.zeros((1, 4, 4, 21), np.float32)
    l = np.full((1, 4, 4), 255, np.int32)
    with pytest.raises(ValueError):
        oracle.e_step(f, l)


def test_oracle_softmax_ce_against_numpy(oracle):
    rng = np.random.default_rng(5)
    z = rng.standard_normal((50, 21)).astype(np.float32) * 4
    y = rng.integers(0, 21, 50).astype(np.uint8)
    y[::7] = 255
    loss, kept, grad = oracle.softmax_ce(z, y)
    keep = y != 255
    zz = z.astype(np.float64)
    lse = np.log(np.exp(zz - zz.max(1, keepdims=True)).sum(1)) + zz.max(1)
    want = (lse - zz[np.arange(50), np.minimum(y, 20)])[keep].mean()
    assert kept == keep.sum()
    assert abs(loss - want) < 1e-12
    sm = np.exp(zz - lse[:, None])
    sm[np.arange(50), np.minimum(y, 20)] -= 1
    sm[~keep] = 0
    assert np.allclose(grad, sm / keep.sum(), atol=1e-15)


def test_oracle_resize_bilinear_against_a_scalar_restatement():
    """network.py:40: the vectorised restatement of TensorFlow's resize_bilinear kernel against the published
    per-pixel formulas written out as scalar float32 loops; plus the properties that need no reference:
    same size = identity, a constant stays constant, the first row/column are copied (lerp 0)."""
    from oracle.oracle import resize_bilinear
    rng = np.random.default_rng(3)
    x = (rng.standard_normal((2, 5, 7, 3)) * 3.0).astype(np.float32)
    H, W = 13, 11
    got = resize_bilinear(x, H, W)
    f32 = np.float32
    sy, sx = f32(5) / f32(H), f32(7) / f32(W)
    for b in range(2):
        for y in range(H):
            fy = f32(y) * sy
            y0 = int(fy); y1 = min(y0 + 1, 4); ly = f32(fy - f32(y0))
            for xo in range(W):
                fx = f32(xo) * sx
                x0 = int(fx); x1 = min(x0 + 1, 6); lx = f32(fx - f32(x0))
                for c in range(3):
                    tl, tr, bl, br = x[b, y0, x0, c], x[b, y0, x1, c], x[b, y1, x0, c], x[b, y1, x1, c]
                    top = f32(tl + f32(f32(tr - tl) * lx))
                    bot = f32(bl + f32(f32(br - bl) * lx))
                    want = f32(top + f32(f32(bot - top) * ly))
                    assert got[b, y, xo, c].view(np.uint32) == want.view(np.uint32)
    assert np.array_equal(resize_bilinear(x, 5, 7), x)
    assert np.array_equal(resize_bilinear(np.full((1, 3, 3, 2), 1.25, np.float32), 9, 8), np.full((1, 9, 8, 2), 1.25, np.float32))
    assert np.array_equal(got[:, 0, 0], x[:, 0, 0])
