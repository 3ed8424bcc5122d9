"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY (ctypes front-end of the checker).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product package never does.

Two checkers live here:
  * ``Oracle``    -- our C restatement (oracle/estep_oracle.c) with explicit RNG state;
  * ``Reference`` -- the reference's own weaklabel.cpp, compiled unmodified into
                     oracle/_ref/ by oracle/Makefile (or the binary it ships), driven
                     exactly as the reference's estep.py:29-31 drives it, one image per
                     call because the reference crashes for batch > 1
                     (weaklabel.cpp:117-118).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(HERE, "libestep_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libweaklabel_ref.so")
SHIPPED_SO = os.path.join(HERE, "_ref", "libweaklabel_shipped.so")

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)


def build(force=False):
    """Compile the checker (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(_ORACLE_SO) or \
            os.path.getmtime(_ORACLE_SO) < os.path.getmtime(os.path.join(HERE, "estep_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, _ORACLE_SO])
    if not os.path.exists(REF_SO):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def resize_nearest(gt, H, W):
    """deeplab.py:110 -- tf.image.resize_nearest_neighbor(label, fc8 size), align_corners=False, restated
    from TensorFlow's documented kernel (third-party to the reference, version unpinned: parity unpinned):
    source row = min(floor(row * float32(Hg / H)), Hg - 1), columns alike.  gt: [B,Hg,Wg] -> [B,H,W]."""
    gt = np.asarray(gt)
    Hg, Wg = gt.shape[1], gt.shape[2]
    sh, sw = np.float32(Hg) / np.float32(H), np.float32(Wg) / np.float32(W)
    ys = np.minimum(np.floor(np.arange(H, dtype=np.float32) * sh).astype(np.int64), Hg - 1)
    xs = np.minimum(np.floor(np.arange(W, dtype=np.float32) * sw).astype(np.int64), Wg - 1)
    return gt[:, ys][:, :, xs]


def resize_bilinear(x, H, W):
    """network.py:40 -- tf.image.resize_bilinear(output, (h, w)) with the TF1 defaults (align_corners=False, no
    half-pixel centres), restated from TensorFlow's published CPU kernel (tensorflow/core/kernels/
    resize_bilinear_op.cc; third-party to the reference, version unpinned: parity unpinned): per output index
    in = index * float32(in_size / out_size), lower = int(in), upper = min(lower + 1, in_size - 1), lerp = in - lower;
    value = top + (bottom - top) * y_lerp with top = tl + (tr - tl) * x_lerp, every step in float32.
    x: [B,h,w,C] float32 -> [B,H,W,C] float32."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    h, w = x.shape[1], x.shape[2]
    sy, sx = np.float32(h) / np.float32(H), np.float32(w) / np.float32(W)
    fy = (np.arange(H, dtype=np.float32) * sy).astype(np.float32)
    fx = (np.arange(W, dtype=np.float32) * sx).astype(np.float32)
    y0, x0 = fy.astype(np.int64), fx.astype(np.int64)
    y1, x1 = np.minimum(y0 + 1, h - 1), np.minimum(x0 + 1, w - 1)
    ly = (fy - y0.astype(np.float32)).astype(np.float32)[None, :, None, None]
    lx = (fx - x0.astype(np.float32)).astype(np.float32)[None, None, :, None]
    tl, tr = x[:, y0][:, :, x0], x[:, y0][:, :, x1]
    bl, br = x[:, y1][:, :, x0], x[:, y1][:, :, x1]
    top = (tl + ((tr - tl).astype(np.float32) * lx).astype(np.float32)).astype(np.float32)
    bot = (bl + ((br - bl).astype(np.float32) * lx).astype(np.float32)).astype(np.float32)
    return (top + ((bot - top).astype(np.float32) * ly).astype(np.float32)).astype(np.float32)


class Oracle:
    """C restatement; batched = B consecutive single-image calls on one RNG stream."""

    def __init__(self, seed=1):
        build()
        self.lib = C.CDLL(_ORACLE_SO)
        L = self.lib
        L.oracle_rng_sizeof.restype = C.c_size_t
        L.oracle_rng_draws.restype = C.c_longlong
        L.oracle_rng_draws.argtypes = [C.c_void_p]
        L.oracle_srand.argtypes = [C.c_void_p, C.c_uint]
        L.oracle_rand.argtypes = [C.c_void_p]
        L.oracle_rand.restype = C.c_int
        L.oracle_rank.argtypes = [C.c_float, C.c_int]
        L.oracle_rank.restype = C.c_int
        L.oracle_e_step.argtypes = [_f32p, _i32p, _i32p, _i32p, C.c_int, C.c_int, C.c_float,
                                    C.c_float, C.c_float, C.c_void_p,
                                    _i32p, _f32p, _f32p, C.POINTER(C.c_double)]
        L.oracle_e_step.restype = C.c_int
        L.oracle_label_set.argtypes = [_i32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p]
        L.oracle_argmax_labels.argtypes = [_f32p, _i32p, _i32p, C.POINTER(C.c_ubyte)]
        L.oracle_softmax_ce.argtypes = [_f32p, C.POINTER(C.c_ubyte), C.c_long, C.c_int, C.c_int,
                                        C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.oracle_softmax_ce.restype = C.c_long
        self._rng = C.create_string_buffer(L.oracle_rng_sizeof())
        self.seed(seed)

    # -- RNG ---------------------------------------------------------------
    def seed(self, seed=1):
        self.lib.oracle_srand(self._rng, seed)

    def rand(self):
        return self.lib.oracle_rand(self._rng)

    @property
    def draws(self):
        return self.lib.oracle_rng_draws(self._rng)

    # -- E-step ------------------------------------------------------------
    def e_step(self, fmap, label, suppress_others=True, num_iter=5, margin_others=1e-5,
               bg=0.4, fg=0.2, probes=False):
        """fmap float32 [B,H,W,C] (any strides the reference accepts), label [B,H,W].
        Returns the shifted map (a new array with fmap's strides untouched is NOT made:
        like the reference, works on a float32 copy made by astype)."""
        f = np.array(fmap, dtype=np.float32, order="C", copy=True)
        l = np.ascontiguousarray(label, dtype=np.int32)
        out = self.e_step_inplace(f, l, suppress_others, num_iter, margin_others, bg, fg, probes)
        return (f,) + out if probes else f

    def e_step_inplace(self, f, l, suppress_others=True, num_iter=5, margin_others=1e-5,
                       bg=0.4, fg=0.2, probes=False):
        assert f.dtype == np.float32 and l.dtype == np.int32 and f.ndim == 4
        B, H, W, Cc = f.shape
        order = np.full((B, num_iter, Cc), -1, np.int32)
        bias = np.zeros((B, num_iter, Cc), np.float32)
        m0 = np.zeros(B, np.float32)
        m1 = np.zeros(B, np.float64)
        rc = self.lib.oracle_e_step(
            f.ctypes.data_as(_f32p), f.ctypes.shape_as(C.c_int), f.ctypes.strides_as(C.c_int),
            l.ctypes.data_as(_i32p), int(bool(suppress_others)), int(num_iter),
            float(margin_others), float(bg), float(fg), self._rng,
            order.ctypes.data_as(_i32p), bias.ctypes.data_as(_f32p),
            m0.ctypes.data_as(_f32p), m1.ctypes.data_as(C.POINTER(C.c_double)))
        if rc != 0:
            raise ValueError("oracle_e_step failed rc=%d (empty label set or bad rank)" % rc)
        return (order, bias, m0, m1) if probes else ()

    def label_set(self, label, num_classes):
        l = np.ascontiguousarray(label, dtype=np.int32)
        B, H, W = l.shape
        out = np.zeros((B, num_classes), np.int32)
        self.lib.oracle_label_set(l.ctypes.data_as(_i32p), B, H, W, num_classes,
                                  out.ctypes.data_as(_i32p))
        return out

    def argmax_labels(self, fmap):
        f = np.asarray(fmap)
        assert f.dtype == np.float32
        out = np.zeros(f.shape[:3], np.uint8)
        self.lib.oracle_argmax_labels(f.ctypes.data_as(_f32p), f.ctypes.shape_as(C.c_int),
                                      f.ctypes.strides_as(C.c_int),
                                      out.ctypes.data_as(C.POINTER(C.c_ubyte)))
        return out

    def softmax_ce(self, logits, labels, ignore_label=255, want_grad=True):
        z = np.ascontiguousarray(logits, dtype=np.float32)
        Cc = z.shape[-1]
        z2 = z.reshape(-1, Cc)
        y = np.ascontiguousarray(labels, dtype=np.uint8).reshape(-1)
        loss = C.c_double(0.0)
        grad = np.zeros(z2.shape, np.float64) if want_grad else None
        kept = self.lib.oracle_softmax_ce(
            z2.ctypes.data_as(_f32p), y.ctypes.data_as(C.POINTER(C.c_ubyte)), z2.shape[0], Cc,
            int(ignore_label), C.byref(loss),
            grad.ctypes.data_as(C.POINTER(C.c_double)) if want_grad else None)
        return loss.value, kept, (grad.reshape(z.shape) if want_grad else None)


class Reference:
    """The reference's weaklabel.cpp (unmodified), called as estep.py:29-31 calls it,
    one image per call.  Uses the process-global libc rand() -- run golden cases in a
    fresh process."""

    def __init__(self, which="ref"):
        path = {"ref": REF_SO, "shipped": SHIPPED_SO}[which]
        if not os.path.exists(path):
            build()
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.path = path
        self.lib = C.cdll.LoadLibrary(path)
        self.lib.e_step.argtypes = [_f32p, _i32p, _i32p, _i32p, C.c_bool, C.c_int,
                                    C.c_float, C.c_float, C.c_float]

    @staticmethod
    def available(which="ref"):
        return os.path.exists({"ref": REF_SO, "shipped": SHIPPED_SO}[which])

    def e_step_image_inplace(self, f1, l1, suppress_others=True, num_iter=5,
                             margin_others=1e-5, bg=0.4, fg=0.2):
        """f1: float32 [1,H,W,C] view (strides honoured), l1: int32 [1,H,W] contiguous."""
        assert f1.shape[0] == 1 and f1.dtype == np.float32 and l1.dtype == np.int32
        self.lib.e_step(f1.ctypes.data_as(_f32p), f1.ctypes.shape_as(C.c_int),
                        f1.ctypes.strides_as(C.c_int), l1.ctypes.data_as(_i32p),
                        bool(suppress_others), int(num_iter), float(margin_others),
                        float(bg), float(fg))

    def e_step(self, fmap, label, **kw):
        f = np.array(fmap, dtype=np.float32, order="C", copy=True)
        l = np.ascontiguousarray(label, dtype=np.int32)
        for b in range(f.shape[0]):
            self.e_step_image_inplace(f[b:b + 1], l[b:b + 1], **kw)
        return f
