"""Device-pointer front end: EStepPlan / OrderStream over the plan API of
include/weaklabel_cuda.h.  torch is used for device memory and streams only (tensor.data_ptr(),
torch.cuda.current_stream()); all compute is in libweaklabel_cuda.so."""
import ctypes as C

import numpy as np

from . import _lib


class OrderStream:
    """Explicit class-order stream == glibc rand() from `seed` (weaklabel.cpp:83)."""

    def __init__(self, seed=1):
        self._L = _lib.load()
        self.handle = C.c_void_p(self._L.weaklabel_rng_create(seed))

    def next(self):
        return self._L.weaklabel_rng_next(self.handle)

    def skip(self, n):
        self._L.weaklabel_rng_skip(self.handle, int(n))

    @property
    def draws(self):
        return self._L.weaklabel_rng_draws(self.handle)

    def __del__(self):
        try:
            self._L.weaklabel_rng_destroy(self.handle)
        except Exception:
            pass


def label_set(label, num_classes):
    """weaklabel.cpp:19-31 on a host int32 [B,H,W] ground-truth map -> uint8 [B,C]."""
    L = _lib.load()
    l = np.ascontiguousarray(label, dtype=np.int32)
    B, H, W = l.shape
    out = np.zeros((B, num_classes), np.uint8)
    _lib.check(L.weaklabel_label_set_host(l.ctypes.data_as(C.POINTER(C.c_int)), B, H, W, num_classes,
                                          out.ctypes.data_as(C.POINTER(C.c_ubyte))))
    return out


def draws_for(n_present, num_iter):
    return _lib.load().weaklabel_draws_for(int(n_present), int(num_iter))


class EStepPlan:
    """One shape on one device.  prepare() fixes label sets + scalars + class order;
    labels()/finalize_map() launch on the current torch stream and never synchronise."""

    def __init__(self, B, H, W, C_, device=0):
        self._L = _lib.load()
        self.B, self.H, self.W, self.C, self.device = B, H, W, C_, device
        h = C.c_void_p()
        _lib.check(self._L.weaklabel_plan_create(C.byref(h), B, H, W, C_, device))
        self.handle = h
        self.num_iter = None

    def tune(self, cluster_size=0, tile_pixels=0, max_clusters=0):
        _lib.check(self._L.weaklabel_plan_tune(self.handle, cluster_size, tile_pixels, max_clusters))
        return self

    def set_path(self, path):
        """"auto" (0), "generic" (1: estep_kernel.cu) or "fast" (2: the on-chip kernel, estep_onchip.cu; labels()
        fails if it is forced and does not apply)."""
        _lib.check(self._L.weaklabel_plan_set_path(self.handle, {"auto": 0, "generic": 1, "fast": 2}.get(path, path)))
        return self

    def profile(self, enable=True, read=False):
        """Enable/disable per-image phase stamps; read=True returns the last run's [B,32] uint64 table."""
        out = np.zeros((self.B, 32), np.uint64) if read else None
        _lib.check(self._L.weaklabel_plan_profile(self.handle, int(bool(enable)),
                                                  out.ctypes.data_as(C.c_void_p) if read else None))
        return out

    @property
    def last_path(self):
        return {0: None, 1: "generic", 2: "fast"}[self._L.weaklabel_plan_last_path(self.handle)]

    def prepare(self, present, suppress_others=True, num_iter=5, margin_others=1e-5, bg=0.4, fg=0.2,
                rng=None):
        """Label sets [B,C] (0/1) + the five scalars of weaklabel.cpp:7 + the class order (drawn here, on
        the host, from `rng` or the process-global rand()).  The tables go up asynchronously on the
        current torch stream; nothing blocks."""
        p = np.ascontiguousarray(present, dtype=np.uint8)
        assert p.shape == (self.B, self.C)
        _lib.check(self._L.weaklabel_plan_prepare_async(
            self.handle, p.ctypes.data_as(C.POINTER(C.c_ubyte)), bool(suppress_others), int(num_iter),
            float(margin_others), float(bg), float(fg), rng.handle if rng is not None else None, self._stream()))
        self.num_iter = int(num_iter)
        return self

    def prepare_from_gt(self, gt, suppress_others=True, num_iter=5, margin_others=1e-5, bg=0.4, fg=0.2, rng=None):
        """deeplab.py:110 + weaklabel.cpp:19-31 on the device: `gt` is the dense ground-truth map, a CUDA
        uint8 / int32 tensor [B,Hg,Wg] at ANY resolution; it is shrunk to the score size by nearest
        neighbour and scanned at stride 2 for the image-level label sets (one small stream sync: the class
        order is drawn on the host from rand())."""
        import torch
        assert gt.is_cuda and gt.is_contiguous() and gt.dim() == 3 and gt.shape[0] == self.B
        assert gt.dtype in (torch.uint8, torch.int32)
        _lib.check(self._L.weaklabel_plan_prepare_device_gt(
            self.handle, gt.data_ptr(), int(gt.dtype == torch.int32), int(gt.shape[1]), int(gt.shape[2]),
            bool(suppress_others), int(num_iter), float(margin_others), float(bg), float(fg),
            rng.handle if rng is not None else None, self._stream()))
        self.num_iter = int(num_iter)
        return self

    def set_outputs(self, sums=True):
        """sums=True (default): labels() also leaves mean_max_0 / mean_max_1 / shift of every image (probes(),
        finalize_map()).  sums=False: labels only -- same labels, bit for bit, but the two order-exact sums are
        formed only for images in which a pixel's argmax can depend on the rounding of the shift."""
        _lib.check(self._L.weaklabel_plan_set_outputs(self.handle, int(bool(sums))))
        return self

    def status(self):
        """Device-side status of the runs since the last call (0 = fine); raises WeakLabelError when a kernel
        reported NaN / Inf scores or a lost arrival.  Synchronises the plan's last stream."""
        code = C.c_int(0)
        _lib.check(self._L.weaklabel_plan_status(self.handle, C.byref(code)))
        return code.value

    def _stream(self):
        """The current torch stream of THE PLAN'S device (a process may drive several GPUs: the current device
        can be another one), else the legacy default stream."""
        try:
            import torch
            if torch.cuda.is_available():
                return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        except ImportError:
            pass
        return C.c_void_p(0)

    def labels(self, fmap, out=None):
        """fmap: CUDA float32 tensor [B,H,W,C] (stride(3)==1, rows contiguous).  -> uint8 [B,H,W]."""
        import torch
        assert fmap.is_cuda and fmap.dtype == torch.float32 and tuple(fmap.shape) == (self.B, self.H, self.W, self.C)
        assert fmap.stride(3) == 1 and fmap.stride(1) == self.W * fmap.stride(2)
        if out is None:
            out = torch.empty((self.B, self.H, self.W), dtype=torch.uint8, device=fmap.device)
        _lib.check(self._L.weaklabel_estep_labels_device(self.handle, fmap.data_ptr(), fmap.stride(0),
                                                         fmap.stride(2), out.data_ptr(), self._stream()))
        return out

    def labels_upsampled(self, fc8, out=None):
        """E-step labels on tf.image.resize_bilinear(fc8, (H, W)) (network.py:40) WITHOUT materialising the
        up-sampled scores: fc8 is the contiguous low-resolution CUDA float32 tensor [B,h,w,C]; the plan was created
        and prepared for the full-resolution shape.  -> uint8 [B,H,W], bit-identical to
        labels(upsample_bilinear(fc8, H, W))."""
        import torch
        assert fc8.is_cuda and fc8.dtype == torch.float32 and fc8.dim() == 4 and fc8.is_contiguous()
        assert fc8.shape[0] == self.B and fc8.shape[3] == self.C
        if out is None:
            out = torch.empty((self.B, self.H, self.W), dtype=torch.uint8, device=fc8.device)
        _lib.check(self._L.weaklabel_estep_labels_upsampled_device(self.handle, fc8.data_ptr(), int(fc8.shape[1]),
                                                                   int(fc8.shape[2]), out.data_ptr(), self._stream()))
        return out

    def labels_dlpack(self, scores, labels):
        """The DLPack front door (weaklabel_estep_labels_dlpack): `scores` / `labels` are ANY objects with
        __dlpack__ (torch, cupy, jax arrays ...) holding float32 [B,H,W,C] / uint8 [B,H,W] CUDA memory.
        Launches on the current torch stream when torch is there, else on the legacy default stream."""
        caps = [scores.__dlpack__(), labels.__dlpack__()]
        get = C.pythonapi.PyCapsule_GetPointer
        get.restype = C.c_void_p
        get.argtypes = [C.py_object, C.c_char_p]
        _lib.check(self._L.weaklabel_estep_labels_dlpack(self.handle, get(caps[0], b"dltensor"), get(caps[1], b"dltensor"),
                                                         self._stream()))
        return labels

    def labels_and_loss(self, fmap, labels=None, want_grad=True):
        """deeplab.py:120-123 + 182 in one call (weaklabel_estep_ce_device): the E-step labels of `fmap`
        (contiguous CUDA float32 [B,H,W,C], the fc8 logits) and the mean softmax cross-entropy of the same
        logits against them, plus its gradient.  -> (labels, loss 0-d, valid 0-d int32, dlogits or None)."""
        import torch
        assert fmap.is_cuda and fmap.dtype == torch.float32 and fmap.is_contiguous()
        assert tuple(fmap.shape) == (self.B, self.H, self.W, self.C)
        if labels is None:
            labels = torch.empty((self.B, self.H, self.W), dtype=torch.uint8, device=fmap.device)
        loss = torch.empty((), dtype=torch.float32, device=fmap.device)
        valid = torch.empty((), dtype=torch.int32, device=fmap.device)
        grad = torch.empty_like(fmap) if want_grad else None
        work = torch.empty(2, dtype=torch.float64, device=fmap.device)
        _lib.check(self._L.weaklabel_estep_ce_device(self.handle, fmap.data_ptr(), labels.data_ptr(), loss.data_ptr(),
                                                     valid.data_ptr(), grad.data_ptr() if want_grad else None,
                                                     work.data_ptr(), self._stream()))
        return labels, loss, valid, grad

    def finalize_map(self, fmap):
        """In place: fmap becomes what the reference's e_step leaves in `map`."""
        _lib.check(self._L.weaklabel_estep_finalize_map_device(self.handle, fmap.data_ptr(), fmap.stride(0),
                                                               fmap.stride(2), self._stream()))
        return fmap

    def probes(self):
        J = self.num_iter
        order = np.full((self.B, J, self.C), -1, np.int32)
        bias = np.zeros((self.B, J, self.C), np.float32)
        m0 = np.zeros(self.B, np.float32)
        m1 = np.zeros(self.B, np.float64)
        _lib.check(self._L.weaklabel_plan_probes(
            self.handle, order.ctypes.data_as(C.POINTER(C.c_int)), bias.ctypes.data_as(C.POINTER(C.c_float)),
            m0.ctypes.data_as(C.POINTER(C.c_float)), m1.ctypes.data_as(C.POINTER(C.c_double))))
        return order, bias, m0, m1

    @property
    def launches(self):
        return self._L.weaklabel_plan_launches(self.handle)

    def close(self):
        if self.handle:
            self._L.weaklabel_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def label_set_device(gt, H, W, num_classes):
    """Image-level label sets of a dense CUDA ground-truth map [B,Hg,Wg] (uint8 / int32) shrunk to HxW
    (deeplab.py:110, weaklabel.cpp:19-31) -> int32 CUDA tensor [B] of class bit masks.  Asynchronous."""
    import torch
    L = _lib.load()
    assert gt.is_cuda and gt.is_contiguous() and gt.dim() == 3 and gt.dtype in (torch.uint8, torch.int32)
    masks = torch.empty(gt.shape[0], dtype=torch.int32, device=gt.device)
    flags = torch.empty(1, dtype=torch.int32, device=gt.device)
    with torch.cuda.device(gt.device):
        _lib.check(L.weaklabel_label_set_device(gt.data_ptr(), int(gt.dtype == torch.int32), int(gt.shape[0]), int(gt.shape[1]),
                                                int(gt.shape[2]), int(H), int(W), int(num_classes), masks.data_ptr(),
                                                flags.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return masks


def fixed_bias_labels(fmap, masks, bias_bg, bias_fg, strong=None):
    """EM-Fixed (the paper's fixed-bias E-step; not in the reference): label = first maximum over the present
    classes of score + bias_bg (class 0) / bias_fg, overridden by `strong` (uint8 [B,H,W], 255 = none) where it
    names a class.  fmap CUDA float32 [B,H,W,C] contiguous, masks int32 CUDA [B] (label_set_device)."""
    import torch
    L = _lib.load()
    assert fmap.is_cuda and fmap.dtype == torch.float32 and fmap.is_contiguous() and fmap.dim() == 4
    B, H, W, Cc = fmap.shape
    assert masks.is_cuda and masks.dtype == torch.int32 and masks.numel() == B
    out = torch.empty((B, H, W), dtype=torch.uint8, device=fmap.device)
    if strong is not None:
        assert strong.is_cuda and strong.dtype == torch.uint8 and strong.is_contiguous() and strong.numel() == B * H * W
    with torch.cuda.device(fmap.device):
        _lib.check(L.weaklabel_fixed_bias_labels_device(fmap.data_ptr(), B, H, W, Cc, masks.data_ptr(), float(bias_bg), float(bias_fg),
                                                        strong.data_ptr() if strong is not None else None, out.data_ptr(),
                                                        C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def upsample_bilinear(fc8, H, W, out=None):
    """tf.image.resize_bilinear(fc8, (H, W)) with the TF1 defaults (network.py:40), materialised on the device:
    fc8 CUDA float32 [B,h,w,C] contiguous -> [B,H,W,C]."""
    import torch
    assert fc8.is_cuda and fc8.dtype == torch.float32 and fc8.dim() == 4 and fc8.is_contiguous()
    B, h, w, Cc = (int(v) for v in fc8.shape)
    if out is None:
        out = torch.empty((B, int(H), int(W), Cc), dtype=torch.float32, device=fc8.device)
    with torch.cuda.device(fc8.device):
        _lib.check(_lib.load().weaklabel_upsample_bilinear_device(fc8.data_ptr(), B, h, w, Cc, int(H), int(W), out.data_ptr(),
                                                                  torch.cuda.current_stream().cuda_stream))
    return out


def softmax_ce(logits, labels, ignore_label=255, want_grad=True):
    """Fused softmax-CE forward(+backward) with ignore_label (deeplab.py:182, network.py:77-89).
    logits CUDA float32 [..., C] contiguous, labels CUDA uint8 [...].
    -> (loss_mean 0-d tensor, valid 0-d int32 tensor, dlogits or None)."""
    import torch
    L = _lib.load()
    assert logits.is_cuda and logits.dtype == torch.float32 and logits.is_contiguous()
    assert labels.is_cuda and labels.dtype == torch.uint8 and labels.is_contiguous()
    Cc = logits.shape[-1]
    n = logits.numel() // Cc
    assert labels.numel() == n
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    valid = torch.empty((), dtype=torch.int32, device=logits.device)
    grad = torch.empty_like(logits) if want_grad else None
    work = torch.empty(2, dtype=torch.float64, device=logits.device)
    with torch.cuda.device(logits.device):          # the launch goes to the tensors' device, whatever the current one is
        _lib.check(L.weaklabel_softmax_ce_device(logits.data_ptr(), labels.data_ptr(), n, Cc, int(ignore_label),
                                                 loss.data_ptr(), valid.data_ptr(),
                                                 grad.data_ptr() if want_grad else None, work.data_ptr(),
                                                 C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return loss, valid, grad
