"""-m gpu: the rest of the C-ABI of libweaklabel_cuda.so around the E-step kernel -- asynchronous prepare,
status word, device label sets (GT shrink), the E-step + loss sequence, the DLPack door, the torch
autograd Function, the pipelined legacy host path (pinned and pageable callers, many blocks) and the
drop-in aliases called from a reference-style directory layout.  Oracle = oracle/ (checker only)."""
import ctypes
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

import cases
import em_adapt_tensorflow_b200 as wl
from oracle.oracle import Oracle, resize_nearest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torch():
    import torch
    return torch


def _oracle(f, l, p):
    orc = Oracle(1)
    fo = np.array(f, dtype=np.float32, order="C", copy=True)
    orc.e_step_inplace(fo, l, **p)
    return fo, orc.argmax_labels(fo)


def test_prepare_is_asynchronous_and_ordered_across_streams():
    """prepare() on one stream, labels() on another, prepare() again while the first run may still be in
    flight: the table upload and the kernels are ordered by events, never by a host synchronisation."""
    torch = _torch()
    f, l, p = cases.build_case("c2_41_b30")
    _, want = _oracle(f, l, p)
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    present = wl.label_set(l, 21)
    plan = wl.EStepPlan(*f.shape)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = []
    for rep in range(6):
        with torch.cuda.stream(s1 if rep % 2 else s2):
            plan.prepare(present, rng=wl.OrderStream(1))
        with torch.cuda.stream(s2 if rep % 2 else s1):
            outs.append(plan.labels(t))
    torch.cuda.synchronize()
    for o in outs:
        assert np.array_equal(o.cpu().numpy(), want)
    assert plan.status() == 0
    plan.close()


def test_status_reports_nan_scores():
    torch = _torch()
    f, l, p = cases.build_case("c1_41_p2")
    f = np.ascontiguousarray(f).copy()
    f[0, 3, 4, :] = np.nan
    plan = wl.EStepPlan(*f.shape)
    plan.prepare(wl.label_set(l, 21), rng=wl.OrderStream(1))
    plan.labels(torch.from_numpy(f).cuda())
    try:
        code = plan.status()
    except wl.WeakLabelError as e:
        assert "inconsistent" in str(e)
    else:
        assert code == 0          # a NaN that never reaches a selection is not an error
    assert plan.status() == 0     # read-and-clear
    plan.close()


@pytest.mark.parametrize("dtype", ["uint8", "int32"])
@pytest.mark.parametrize("name,Hg,Wg", [("c2_41_b30", 321, 321), ("c2_41_b30", 41, 41), ("odd_only", 0, 0),
                                        ("void_labels", 0, 0), ("p5_seq3", 100, 77)])
def test_device_label_sets_and_gt_shrink(name, Hg, Wg, dtype, oracle):
    """deeplab.py:110 + weaklabel.cpp:19-31 on the device against the oracle's label_set of the NumPy shrink."""
    torch = _torch()
    f, l, p = cases.build_case(name)
    B, H, W, C = f.shape
    if Hg:
        rng = np.random.default_rng(5)
        sets = [np.flatnonzero(row) for row in wl.label_set(l, C)]
        gt = np.stack([cases.paint_label_map(Hg, Wg, list(s), rng) for s in sets]).astype(np.int32)
        gt[:, ::7, ::5] = 255                              # void pixels
    else:
        gt = l
        if dtype == "uint8":
            gt = np.where((gt < 0) | (gt > 255), 255, gt)
    shrunk = resize_nearest(gt, H, W).astype(np.int32)
    want = np.asarray(oracle.label_set(np.ascontiguousarray(shrunk), C)).reshape(B, C)
    if want.sum(axis=1).min() == 0:
        pytest.skip("shrink left an empty label set")
    g = torch.from_numpy(np.ascontiguousarray(gt.astype(dtype))).cuda()
    masks = wl.label_set_device(g, H, W, C).cpu().numpy().astype(np.int64) & 0xffffffff
    got = ((masks[:, None] >> np.arange(C)[None, :]) & 1).astype(np.uint8)
    assert np.array_equal(got, want.astype(np.uint8))
    # the whole prepare from the device map gives the labels of the host path
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    a = wl.EStepPlan(B, H, W, C).prepare_from_gt(g, p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                                                 rng=wl.OrderStream(1))
    b = wl.EStepPlan(B, H, W, C).prepare(want, p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                                         rng=wl.OrderStream(1))
    assert torch.equal(a.labels(t), b.labels(t))
    a.close(); b.close()


@pytest.mark.parametrize("name", ["c3_321_b30", "c2_41_b30", "p21_41"])
def test_estep_and_loss_in_one_call(name, oracle):
    """weaklabel_estep_ce_device: labels bit-exact, loss and gradient within 1e-5 relative of the fp64 oracle
    (the tolerance BASELINE.json's north_star states)."""
    torch = _torch()
    f, l, p = cases.build_case(name)
    _, want_lab = _oracle(f, l, p)
    z = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    plan = wl.EStepPlan(*f.shape)
    plan.prepare(wl.label_set(l, f.shape[3]), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                 rng=wl.OrderStream(1))
    lab, loss, valid, grad = plan.labels_and_loss(z)
    torch.cuda.synchronize()
    assert np.array_equal(lab.cpu().numpy(), want_lab)
    want_loss, kept, want_grad = oracle.softmax_ce(np.ascontiguousarray(f).reshape(-1, f.shape[3]), want_lab.reshape(-1))
    assert int(valid.item()) == kept == want_lab.size
    tol = 1e-5
    assert abs(float(loss.item()) - want_loss) <= tol * abs(want_loss)
    g = grad.cpu().numpy().astype(np.float64).reshape(-1, f.shape[3])
    assert np.abs(g - want_grad).max() <= tol * np.abs(want_grad).max()
    assert plan.status() == 0
    plan.close()


@pytest.mark.parametrize("n,C", [(256 * 7, 21), (256 * 3 + 17, 21), (100, 5), (256 * 9, 32), (1, 21), (256 * 4 + 1, 2)])
def test_softmax_ce_tiles_and_tails(n, C, oracle):
    """Full bulk tiles, the ragged last tile, small class counts, a misaligned (sliced) tensor."""
    torch = _torch()
    rng = np.random.default_rng(n + C)
    z = (rng.standard_normal((n + 1, C)) * 4).astype(np.float32)
    y = rng.integers(0, C, n + 1).astype(np.uint8)
    y[rng.random(n + 1) < 0.2] = 255
    for off in (0, 1):                                      # off = 1: data pointer 4 * C bytes past an aligned one
        zz, yy = z[off:off + n], y[off:off + n]
        want_loss, kept, want_grad = oracle.softmax_ce(zz, yy)
        zt = torch.from_numpy(z).cuda()[off:off + n]
        yt = torch.from_numpy(y).cuda()[off:off + n].contiguous()
        loss, valid, grad = wl.softmax_ce(zt, yt)
        assert int(valid.item()) == kept
        if kept:
            assert abs(float(loss.item()) - want_loss) <= 1e-5 * max(abs(want_loss), 1e-30)
            assert np.abs(grad.cpu().numpy().astype(np.float64) - want_grad).max() <= 1e-5 * max(np.abs(want_grad).max(), 1e-30)


def test_dlpack_door():
    torch = _torch()
    f, l, p = cases.build_case("p5_seq3")
    _, want = _oracle(f, l, p)

    class Foreign:                                          # any producer: only __dlpack__ is used
        def __init__(self, t):
            self.t = t

        def __dlpack__(self, stream=None):
            return self.t.__dlpack__()

    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    padded = torch.zeros((f.shape[0], f.shape[1], f.shape[2], 24), device="cuda")
    padded[..., :21] = t
    plan = wl.EStepPlan(*f.shape)
    plan.prepare(wl.label_set(l, 21), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"], rng=wl.OrderStream(1))
    out = torch.empty(f.shape[:3], dtype=torch.uint8, device="cuda")
    plan.labels_dlpack(Foreign(t), Foreign(out))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)
    out.zero_()
    plan.labels_dlpack(Foreign(padded[..., :21]), Foreign(out))     # strided view: pixel stride 24
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)
    with pytest.raises(wl.WeakLabelError):
        plan.labels_dlpack(Foreign(t.double()), Foreign(out))
    with pytest.raises(wl.WeakLabelError):
        plan.labels_dlpack(Foreign(t.cpu()), Foreign(out))
    plan.close()


def test_autograd_function_matches_torch_cross_entropy_on_oracle_labels():
    """EMAdaptLoss (deeplab.py:109-124 + 180-186 as one Function): value and gradient equal
    torch.autograd through F.cross_entropy on the ORACLE's labels within 1e-5."""
    torch = _torch()
    from em_adapt_tensorflow_b200.loss import EMAdaptLoss
    f, l, p = cases.build_case("c2_41_b30")
    _, want_lab = _oracle(f, l, p)
    crit = EMAdaptLoss(bg_p=p["bg"], fg_p=p["fg"], num_iter=p["num_iter"], suppress_others=p["suppress_others"],
                       margin_others=p["margin_others"], seed=1)
    fc8 = torch.from_numpy(np.ascontiguousarray(f)).cuda().requires_grad_(True)
    gt = torch.from_numpy(np.ascontiguousarray(l)).cuda()
    loss, labels = crit(fc8 * 1.0, gt)
    (loss * 3.0).backward()
    assert np.array_equal(labels.cpu().numpy(), want_lab)
    ref_in = torch.from_numpy(np.ascontiguousarray(f)).cuda().double().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(ref_in.reshape(-1, 21), torch.from_numpy(want_lab).cuda().reshape(-1).long())
    (ref * 3.0).backward()
    assert abs(float(loss.item()) - float(ref.item())) <= 1e-5 * float(ref.item())
    assert (fc8.grad.double() - ref_in.grad).abs().max().item() <= 1e-5 * ref_in.grad.abs().max().item()


@pytest.mark.parametrize("pinned", [False, True])
@pytest.mark.parametrize("name,block_mb", [("c2_41_b30", 1), ("c3_321_b30", 20), ("stride_padbatch", 0), ("c4_513_b4", 30)])
def test_legacy_pipeline_blocks(name, block_mb, pinned, golden):
    """e_step on host buffers cut into many pipeline blocks (WEAKLABEL_BLOCK_MB), pageable (numpy) and pinned
    callers: same bits as the reference's own outputs (golden hashes), rand() consumed in image order."""
    torch = _torch()
    meta, _ = golden
    f, l, p = cases.build_case(name)
    code = r"""
import ctypes, sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
import cases, torch
import em_adapt_tensorflow_b200 as wl
f, l, p = cases.build_case(%r)
if %r:      # the same (possibly padded) layout inside page-locked memory
    base = f
    while getattr(base, "base", None) is not None:
        base = base.base
    base = np.asarray(base)
    hold = torch.empty(base.size, dtype=torch.float32).pin_memory()
    flat = hold.numpy()
    flat[:] = base.reshape(-1)
    off = (f.__array_interface__["data"][0] - base.__array_interface__["data"][0]) // 4
    f = np.lib.stride_tricks.as_strided(flat[off:], f.shape, f.strides)
ctypes.CDLL("libc.so.6").srand(1)
wl.estep_c(f, l, p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"])
print("HASH", cases.sha1(np.ascontiguousarray(f)), cases.sha1(np.argmax(f, axis=3).astype(np.uint8)))
""" % (ROOT, os.path.join(ROOT, "tests"), name, pinned)
    env = dict(os.environ)
    if block_mb:
        env["WEAKLABEL_BLOCK_MB"] = str(block_mb)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [x for x in r.stdout.splitlines() if x.startswith("HASH")][0].split()
    assert line[1] == meta[name]["map_sha1"] and line[2] == meta[name]["labels_sha1"]


def test_drop_in_aliases_called_from_reference_layout(tmp_path, golden):
    """The reference's callers load ./libweaklabel.so (estep.py:9-12), ./estep/libweaklabel.so
    (estep/estep.py:11) ... relative to their own directory: copy the aliases into such a layout and CALL
    e_step through each of them, with the reference's own ctypes prototype."""
    meta, _ = golden
    wl.load()
    drop = os.path.join(os.path.dirname(wl.lib_path()), "dropin")
    work = tmp_path / "refstyle"
    shutil.copytree(drop, work, symlinks=False)
    code = r"""
import ctypes, sys, numpy as np
from ctypes import POINTER, c_float, c_int, c_bool
sys.path.insert(0, %r)
import cases
for rel in ("libweaklabel.so", "libestep.so", "estep/libweaklabel.so", "pythonlib/estep/libweaklabel.so"):
    lib = ctypes.cdll.LoadLibrary("./" + rel)
    fn = lib.e_step
    fn.restype = POINTER(c_float)
    f, l, p = cases.build_case("p5_seq3")
    ctypes.CDLL("libc.so.6").srand(1)
    fn(f.ctypes.data_as(POINTER(c_float)), f.ctypes.shape_as(c_int), f.ctypes.strides_as(c_int),
       l.ctypes.data_as(POINTER(c_int)), c_bool(p["suppress_others"]), c_int(p["num_iter"]), c_float(p["margin_others"]),
       c_float(p["bg"]), c_float(p["fg"]))
    print("HASH", rel, cases.sha1(np.ascontiguousarray(f)))
""" % (os.path.join(ROOT, "tests"),)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(work), timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [x.split() for x in r.stdout.splitlines() if x.startswith("HASH")]
    assert len(lines) == 4
    for _, rel, h in lines:
        assert h == meta["p5_seq3"]["map_sha1"], rel


@pytest.mark.parametrize("name", ["c3_321_b30", "c2_41_b30", "p5_seq3", "shape_100x37"])
@pytest.mark.parametrize("with_strong", [False, True])
def test_fixed_bias_labels_em_fixed(name, with_strong):
    """EM-Fixed (the paper's fixed-bias E-step, readme.md:3; absent from the reference -- parity unpinned):
    against a NumPy restatement: first maximum over the present classes of score + bias, strong labels win."""
    torch = _torch()
    f, l, p = cases.build_case(name)
    f = np.ascontiguousarray(f)
    B, H, W, C = f.shape
    present = wl.label_set(l, C).astype(bool)
    bias = np.where(np.arange(C) == 0, np.float32(0.7), np.float32(-0.3)).astype(np.float32)
    shifted = (f + bias[None, None, None, :]).astype(np.float32)
    shifted[~np.broadcast_to(present[:, None, None, :], f.shape)] = -np.inf
    want = np.argmax(shifted, axis=3).astype(np.uint8)
    strong = None
    if with_strong:
        rng = np.random.default_rng(3)
        strong = np.where(rng.random((B, H, W)) < 0.3, rng.integers(0, C, (B, H, W)), 255).astype(np.uint8)
        want = np.where(strong < C, strong, want).astype(np.uint8)
    t = torch.from_numpy(f).cuda()
    masks = wl.label_set_device(torch.from_numpy(np.ascontiguousarray(l)).cuda(), H, W, C)
    got = wl.fixed_bias_labels(t, masks, 0.7, -0.3, torch.from_numpy(strong).cuda() if with_strong else None)
    assert np.array_equal(got.cpu().numpy(), want)
    if not with_strong:                                   # a misaligned score map takes the plain-load path: same labels
        flat = torch.empty(f.size + 1, dtype=torch.float32, device="cuda")
        flat[1:] = t.reshape(-1)
        assert torch.equal(wl.fixed_bias_labels(flat[1:].view(B, H, W, C), masks, 0.7, -0.3), got)


@pytest.mark.parametrize("name", ["c3_321_b30", "c4_513_b4", "p5_seq3", "iter9", "shape_100x37"])
def test_finalize_map_tiles_equal_the_simple_kernel(name, monkeypatch):
    """The in-place map on bulk tiles (stream_kernels.cu) against the one-thread-per-pixel kernel and the oracle."""
    torch = _torch()
    f, l, p = cases.build_case(name)
    want, _ = _oracle(f, l, p)
    outs = []
    for simple in (False, True):
        if simple:
            monkeypatch.setenv("WEAKLABEL_FINALIZE_SIMPLE", "1")
        t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
        plan = wl.EStepPlan(*f.shape)
        plan.prepare(wl.label_set(l, f.shape[3]), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                     rng=wl.OrderStream(1))
        plan.labels(t)
        plan.finalize_map(t)
        outs.append(t.cpu().numpy())
        plan.close()
    assert outs[0].tobytes() == want.tobytes() and outs[1].tobytes() == want.tobytes()
