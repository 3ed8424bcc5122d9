"""-m gpu: bilinear up-sampling of the fc8 scores (network.py:40) -- the materialising kernel against the oracle's
restatement of TensorFlow's kernel (bit-exact), and the E-step with the up-sampling FUSED into its producers
 against (a) the same kernel on the materialised scores and (b) the oracle's E-step on the
oracle's up-sampled scores: identical labels.  Oracle = oracle/ (checker only)."""
import numpy as np
import pytest

import cases
import em_adapt_tensorflow_b200 as wl
from oracle.oracle import Oracle, resize_bilinear

pytestmark = pytest.mark.gpu
PARAMS = dict(suppress_others=True, num_iter=5, margin_others=1e-5, bg=0.4, fg=0.2)


def _torch():
    import torch
    return torch


@pytest.mark.parametrize("B,h,w,C,H,W", [(2, 41, 41, 21, 321, 321), (1, 65, 65, 21, 513, 513), (3, 5, 7, 4, 13, 11),
                                          (1, 41, 41, 21, 41, 41), (2, 9, 6, 21, 8, 5), (1, 1, 1, 3, 17, 4),
                                          (1, 2, 1300, 21, 3, 1400)])      # the last: a row pair beyond shared memory
def test_upsample_kernel_is_bit_exact_with_the_oracle(B, h, w, C, H, W):
    torch = _torch()
    rng = np.random.default_rng(B * 1000 + h * 10 + H)
    x = (rng.standard_normal((B, h, w, C)) * 3.0).astype(np.float32)
    want = resize_bilinear(x, H, W)
    got = wl.upsample_bilinear(torch.from_numpy(x).cuda(), H, W).cpu().numpy()
    assert got.shape == (B, H, W, C)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    if (h, w) == (H, W):
        assert np.array_equal(got, x)              # same size: every lerp is 0


def _label_maps(sets, H, W, seed):
    rng = np.random.default_rng(seed)
    return np.stack([cases.paint_label_map(H, W, s, rng) for s in sets]).astype(np.int32)


def _check_fused(fc8, sets, H, W, params=PARAMS):
    torch = _torch()
    B, h, w, C = fc8.shape
    labmap = _label_maps(sets, H, W, 5)
    present = wl.label_set(labmap, C)
    up = resize_bilinear(fc8, H, W)
    orc = Oracle(1)
    fo = up.copy()
    orc.e_step_inplace(fo, labmap, **params)
    want = orc.argmax_labels(fo)
    plan = wl.EStepPlan(B, H, W, C)
    plan.prepare(present, rng=wl.OrderStream(1), **params)
    t = torch.from_numpy(fc8).cuda()
    fused = plan.labels_upsampled(t).cpu().numpy()
    assert plan.status() == 0
    plan.prepare(present, rng=wl.OrderStream(1), **params)
    mat = plan.labels(wl.upsample_bilinear(t, H, W)).cpu().numpy()
    assert plan.status() == 0
    plan.close()
    assert np.array_equal(mat, want), "materialised path differs from the oracle in %d pixels" % int((mat != want).sum())
    assert np.array_equal(fused, want), "fused path differs from the oracle in %d pixels" % int((fused != want).sum())


def test_fused_upsampling_config3_shape():
    """BASELINE.json configs[2]: 321 x 321 scores = the 41 x 41 fc8 tensor up-sampled; every kind of label set."""
    rng = np.random.default_rng(77)
    sets = [[0], [15], [0, 7], [0, 3, 12], [0, 2, 9, 20], [5, 6], list(range(21))]
    fc8 = (rng.standard_normal((len(sets), 41, 41, 21)) * 3.0).astype(np.float32)
    _check_fused(fc8, sets, 321, 321)


def test_fused_upsampling_integer_scores_many_exact_ties():
    """Integer fc8 values: neighbouring pixels and classes tie exactly after interpolation -- the paths that look at a
    pixel's original scores again (absent classes at a tie, exact labels) must interpolate the same bits."""
    rng = np.random.default_rng(78)
    sets = [[0, 4], [9], [0, 1, 2], [0, 5, 6, 7]]
    fc8 = rng.integers(-2, 3, size=(len(sets), 11, 11, 21)).astype(np.float32)
    _check_fused(fc8, sets, 88, 88)
    _check_fused(fc8, sets, 83, 61)


def test_fused_upsampling_metric_shape():
    """513 x 513 from 65 x 65 (the shape the metric is quoted on), clusters of 9."""
    rng = np.random.default_rng(79)
    sets = [[0, 11], [0, 3, 8, 14]]
    fc8 = (rng.standard_normal((len(sets), 65, 65, 21)) * 3.0).astype(np.float32)
    _check_fused(fc8, sets, 513, 513)


def test_fused_upsampling_same_size_and_small_ratio_take_the_global_taps():
    """Same size (every lerp 0: the labels of the tensor itself) and a ratio close to 1: the low-resolution rows of a
    tile do not fit the staging ring, the producers interpolate from global memory."""
    torch = _torch()
    rng = np.random.default_rng(80)
    sets = [[0, 6], [0, 1, 19]]
    fc8 = (rng.standard_normal((len(sets), 200, 200, 21)) * 3.0).astype(np.float32)
    _check_fused(fc8, sets, 200, 200)
    _check_fused(fc8, sets, 231, 207)
    labmap = _label_maps(sets, 200, 200, 5)
    plan = wl.EStepPlan(len(sets), 200, 200, 21)
    plan.prepare(wl.label_set(labmap, 21), rng=wl.OrderStream(1), **PARAMS)
    t = torch.from_numpy(fc8).cuda()
    a = plan.labels_upsampled(t).cpu().numpy()
    plan.prepare(wl.label_set(labmap, 21), rng=wl.OrderStream(1), **PARAMS)
    b = plan.labels(t).cpu().numpy()
    plan.close()
    assert np.array_equal(a, b)


def test_fused_upsampling_needs_the_onchip_kernel():
    torch = _torch()
    sets = [[0, 4]]
    labmap = _label_maps(sets, 64, 64, 1)
    plan = wl.EStepPlan(1, 64, 64, 21)
    plan.prepare(wl.label_set(labmap, 21), rng=wl.OrderStream(1), suppress_others=False, num_iter=2, margin_others=1e-5,
                 bg=0.4, fg=0.2)
    with pytest.raises(wl.WeakLabelError):
        plan.labels_upsampled(torch.zeros((1, 8, 8, 21), dtype=torch.float32, device="cuda"))
    plan.close()
