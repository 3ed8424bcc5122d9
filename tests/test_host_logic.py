"""Host-side logic of libweaklabel_cuda.so that needs no device: the explicit class-order
stream, label sets, the order table, sharding, the order-exact summation scheme, and the
"fails loudly without CUDA" contract."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

import cases
import em_adapt_tensorflow_b200 as wl
from em_adapt_tensorflow_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rng_equals_oracle_and_first_draws(oracle):
    for seed in (1, 7, 99991):
        s = wl.OrderStream(seed)
        oracle.seed(seed)
        assert [s.next() for _ in range(500)] == [oracle.rand() for _ in range(500)]
    s = wl.OrderStream(1)
    assert [s.next() for _ in range(3)] == [1804289383, 846930886, 1681692777]   # glibc, seed 1
    oracle.seed(1)


def test_rng_skip_and_draw_count():
    a, b = wl.OrderStream(3), wl.OrderStream(3)
    for _ in range(137):
        a.next()
    b.skip(137)
    assert a.draws == b.draws == 137
    assert a.next() == b.next()


def test_draws_for_matches_reference_consumption():
    # std::random_shuffle on the P-1 tail entries draws max(P-2,0) times (weaklabel.cpp:83)
    assert [wl.draws_for(p, 5) for p in (1, 2, 3, 21)] == [0, 0, 5, 95]
    assert wl.draws_for(4, 0) == 0


@pytest.mark.parametrize("name", ["odd_only", "void_labels", "c2_41_b30", "shape_7x5", "p3_seq4"])
def test_label_set_matches_oracle(name, oracle):
    f, l, p = cases.build_case(name)
    got = wl.label_set(l, 21)
    want = oracle.label_set(l, 21)
    assert np.array_equal(got.astype(np.int32), want)


def test_label_set_odd_coordinates_dropped():
    lab = cases._lab_odd_only(41, 41)
    s = wl.label_set(lab, 21)[0]
    assert s[0] == 1 and s[7] == 0 and s[9] == 1      # weaklabel.cpp:24-25 stride-2 scan


def test_label_set_rejects_negative():
    lab = np.zeros((1, 4, 4), np.int32)
    lab[0, 2, 2] = -3                                  # UB in the reference (weaklabel.cpp:28)
    with pytest.raises(wl.WeakLabelError):
        wl.label_set(lab, 21)


@pytest.mark.parametrize("name", ["p5_seq3", "p3_seq4", "c2_41_b30", "iter9", "p21_41", "p20_nobg_24"])
def test_order_table_matches_oracle(name):
    from oracle.oracle import Oracle
    f, l, p = cases.build_case(name)
    orc = Oracle(1)
    fo = np.array(f, dtype=np.float32, order="C", copy=True)
    order_o, _, _, _ = orc.e_step_inplace(fo, l, probes=True, **p)
    got = wl.draw_order(wl.label_set(l, 21), p["num_iter"], wl.OrderStream(1))
    assert np.array_equal(got, order_o)


def test_order_table_global_rand_equals_explicit_stream():
    """rng == NULL consumes libc rand() exactly like the explicit stream from seed 1."""
    code = (
        "import sys;sys.path.insert(0,%r);sys.path.insert(0,%r)\n"
        "import numpy as np,cases,em_adapt_tensorflow_b200 as wl\n"
        "f,l,p=cases.build_case('p5_seq3')\n"
        "ps=wl.label_set(l,21)\n"
        "a=wl.draw_order(ps,5,None);b=wl.draw_order(ps,5,wl.OrderStream(1))\n"
        "print('OK' if np.array_equal(a,b) else 'BAD')" % (ROOT, os.path.join(ROOT, "tests")))
    assert subprocess.check_output([sys.executable, "-c", code], text=True).strip().endswith("OK")


def test_empty_label_set_is_an_error():
    with pytest.raises(wl.WeakLabelError):
        wl.draw_order(np.zeros((1, 21), np.uint8), 5, wl.OrderStream(1))


def test_shard_ranges_partition():
    for B, W in ((256, 8), (30, 4), (7, 3), (2, 4)):
        r = [wl.shard_range(B, W, k) for k in range(W)]
        assert r[0][0] == 0 and r[-1][1] == B
        assert all(r[k][1] == r[k + 1][0] for k in range(W - 1))
        assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_sharded_order_equals_sequential_order():
    rng = np.random.default_rng(11)
    sets = cases.voc_label_sets(37, rng, forced={3: list(range(21)), 20: [4, 5, 6, 7]})
    present = np.zeros((37, 21), np.uint8)
    for b, s in enumerate(sets):
        present[b, s] = 1
    whole = wl.draw_order(present, 5, wl.OrderStream(1))
    for W in (2, 3, 8):
        parts = []
        for k in range(W):
            lo, hi = wl.shard_range(37, W, k)
            parts.append(wl.draw_order(present[lo:hi], 5, wl.positioned_stream(present, 5, lo)))
        assert np.array_equal(np.concatenate(parts), whole)


def test_seqsum_host_check(tmp_path):
    """csrc/seqsum.cuh (order-exact parallel evaluation of the reference's sequential float /
    double sums, weaklabel.cpp:69-80,102-106) against a plain loop, on the host."""
    exe = str(tmp_path / "seqsum_test")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe,
                           os.path.join(ROOT, "tests", "cpu", "seqsum_test.cpp")])
    r = subprocess.run([exe, "120"], capture_output=True, text=True)
    assert r.returncode == 0 and " bad 0 " in r.stdout and "MISMATCH" not in r.stdout, r.stdout


def test_e_step_without_cuda_aborts_loudly():
    """No CPU fallback: on a box without a CUDA device the legacy entry point must print and
    abort (its signature has no error channel), never return the map untouched."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    code = (
        "import sys;sys.path.insert(0,%r)\n"
        "import numpy as np,em_adapt_tensorflow_b200 as wl\n"
        "f=np.zeros((1,4,4,21),np.float32);l=np.zeros((1,4,4),np.int32)\n"
        "wl.estep_c(f,l,True,5,1e-5,0.4,0.2);print('RETURNED')" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode != 0 and "RETURNED" not in r.stdout
    assert "no CUDA device" in r.stderr and "no CPU fallback" in r.stderr


def test_plan_create_without_cuda_reports_error():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    with pytest.raises(wl.WeakLabelError, match="no CPU fallback"):
        wl.EStepPlan(1, 4, 4, 21)
