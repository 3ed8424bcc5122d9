"""weaklabel_softmax_ce_static: the loss entry point that allocates nothing (include/weaklabel_cuda.h 6b).

No counterpart in the reference.  The check is against the existing entry point
(weaklabel_softmax_ce_device, which takes a caller-allocated scratch buffer) through the C-ABI, on two
calls in a row with DIFFERENT inputs: the static scratch keeps the previous call's contents, so the
second call proves that the entry point clears what must start at zero.  Collected last on purpose
(pytest -x must not hide other tests behind a failure here); runs alone in a fresh
process as well as after the rest of the suite.
"""

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _run(L, fn, z, y, work=None):
    """One call of `fn` (the static or the device entry) on device tensors; returns (loss, valid, grad)."""
    from em_adapt_tensorflow_b200 import _lib
    n, Cc = z.shape
    loss = torch.full((1,), float("nan"), dtype=torch.float32, device=z.device)
    valid = torch.full((1,), -7, dtype=torch.int32, device=z.device)
    grad = torch.full_like(z, float("nan"))
    stream = torch.cuda.current_stream().cuda_stream
    args = [z.data_ptr(), y.data_ptr(), n, Cc, 255, loss.data_ptr(), valid.data_ptr(), grad.data_ptr()]
    if work is not None:
        args.append(work.data_ptr())
    args.append(stream)
    _lib.check(fn(*args))
    torch.cuda.synchronize()
    return float(loss.item()), int(valid.item()), grad


def _case(seed, n, Cc, ignore_frac):
    rng = np.random.default_rng(seed)
    z = (rng.standard_normal((n, Cc)) * 3.0).astype(np.float32)
    y = rng.integers(0, Cc, size=n).astype(np.uint8)
    y[rng.random(n) < ignore_frac] = 255
    return torch.from_numpy(z).cuda(), torch.from_numpy(y).cuda()


def test_static_scratch_entry_matches_the_device_entry_twice_in_a_row():
    from em_adapt_tensorflow_b200 import _lib
    L = _lib.load()
    work = torch.zeros(4, dtype=torch.float64, device="cuda")          # the existing entry's scratch (>= 16 bytes)
    # two calls in a row, different sizes / classes / ignore fractions; 70001 pixels = 273 full tiles + a ragged one
    for seed, n, Cc, ign in ((11, 70001, 21, 0.1), (12, 33000, 21, 0.5), (13, 513, 5, 0.0)):
        z, y = _case(seed, n, Cc, ign)
        want_loss, want_valid, want_grad = _run(L, L.weaklabel_softmax_ce_device, z, y, work)
        got_loss, got_valid, got_grad = _run(L, L.weaklabel_softmax_ce_static, z, y)
        assert got_valid == want_valid == int((y != 255).sum().item())
        # same kernels, same arithmetic per pixel: the gradient is bit-identical; the mean is reduced in double
        # with atomics whose order varies from launch to launch -> the suite's loss tolerance (1e-5 relative)
        assert torch.equal(got_grad, want_grad)
        assert abs(got_loss - want_loss) <= 1e-5 * max(1.0, abs(want_loss))
        assert np.isfinite(got_loss)


def test_static_scratch_entry_refuses_sizes_beyond_its_maximum_and_bad_arguments():
    from em_adapt_tensorflow_b200 import _lib
    L = _lib.load()
    z, y = _case(3, 256, 21, 0.0)
    loss = torch.zeros(1, dtype=torch.float32, device="cuda")
    rc = L.weaklabel_softmax_ce_static(z.data_ptr(), y.data_ptr(), 2 ** 31, 21, 255, loss.data_ptr(), None, None, None)
    assert rc != 0
    assert b"WEAKLABEL_CE_STATIC_MAX_PIXELS" in L.weaklabel_last_error()
    assert L.weaklabel_softmax_ce_static(None, y.data_ptr(), 256, 21, 255, loss.data_ptr(), None, None, None) != 0
    assert L.weaklabel_softmax_ce_static(z.data_ptr(), y.data_ptr(), 256, 33, 255, loss.data_ptr(), None, None, None) != 0
    # and it still works afterwards (forward only: no gradient buffer, no valid output)
    _lib.check(L.weaklabel_softmax_ce_static(z.data_ptr(), y.data_ptr(), 256, 21, 255, loss.data_ptr(), None, None,
                                             torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    want = torch.nn.functional.cross_entropy(z, y.long()).item()
    assert abs(loss.item() - want) <= 1e-5 * max(1.0, abs(want))
