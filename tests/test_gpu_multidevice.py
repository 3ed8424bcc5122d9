"""-m gpu, needs two GPUs in ONE process (skipped otherwise): everything per-device in the library -- function
attributes of the kernels, the static scratch of weaklabel_softmax_ce_static, the plan's stream -- with the
CURRENT device deliberately left at 0 while the work goes to device 1, and both devices used in turn."""
import numpy as np
import pytest

import cases
import em_adapt_tensorflow_b200 as wl
from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu


def _torch2():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    return torch


def test_plan_on_device_1_while_device_0_is_current():
    torch = _torch2()
    torch.cuda.set_device(0)
    for name in ("c2_41_b30", "c3_321_randn"):
        f, l, p = cases.build_case(name)
        orc = Oracle(1)
        fo = np.array(f, dtype=np.float32, order="C", copy=True)
        orc.e_step_inplace(fo, l, **p)
        want = orc.argmax_labels(fo)
        for dev in (1, 0, 1):
            plan = wl.EStepPlan(*f.shape, device=dev)
            plan.prepare(wl.label_set(l, f.shape[3]), rng=wl.OrderStream(1), **p)
            t = torch.from_numpy(np.ascontiguousarray(f)).to("cuda:%d" % dev)
            got = plan.labels(t)
            assert got.device.index == dev
            assert np.array_equal(got.cpu().numpy(), want)
            assert plan.status() == 0
            plan.close()
    assert torch.cuda.current_device() == 0


def test_streaming_kernels_and_static_scratch_on_both_devices():
    torch = _torch2()
    from em_adapt_tensorflow_b200 import _lib
    L = _lib.load()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(5)
    z = (rng.standard_normal((70001, 21)) * 3.0).astype(np.float32)
    y = rng.integers(0, 21, size=70001).astype(np.uint8)
    y[rng.random(70001) < 0.2] = 255
    fc8 = (rng.standard_normal((2, 9, 11, 21)) * 3.0).astype(np.float32)
    from oracle.oracle import resize_bilinear
    want_up = resize_bilinear(fc8, 40, 37)
    for dev in (1, 0, 1):
        d = "cuda:%d" % dev
        zt, yt = torch.from_numpy(z).to(d), torch.from_numpy(y).to(d)
        loss, valid, grad = wl.softmax_ce(zt, yt)
        ref = torch.nn.functional.cross_entropy(zt, yt.long(), ignore_index=255)
        assert abs(loss.item() - ref.item()) <= 1e-5 * max(1.0, abs(ref.item()))
        assert int(valid.item()) == int((y != 255).sum())
        # the entry point without allocation: its scratch is a per-device copy of a static array
        loss2 = torch.full((1,), float("nan"), device=d)
        grad2 = torch.full_like(zt, float("nan"))
        with torch.cuda.device(dev):
            _lib.check(L.weaklabel_softmax_ce_static(zt.data_ptr(), yt.data_ptr(), z.shape[0], 21, 255, loss2.data_ptr(), None,
                                                     grad2.data_ptr(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize(dev)
        assert torch.equal(grad2, grad)
        assert abs(loss2.item() - loss.item()) <= 1e-5 * max(1.0, abs(loss.item()))
        up = wl.upsample_bilinear(torch.from_numpy(fc8).to(d), 40, 37)
        assert up.device.index == dev and np.array_equal(up.cpu().numpy(), want_up)
    assert torch.cuda.current_device() == 0
