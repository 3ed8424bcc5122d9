"""Image sharding of the E-step across ranks (one process per GPU).

The E-step is independent per image (estep/weaklabel.cpp:34: the loop body touches image b
only), so a batch shards by contiguous image blocks with NO data-path collective.  The one
cross-image coupling is the class-order stream: image b's permutation depends on how many
rand() draws the images before it consumed (weaklabel.cpp:83; num_iter * max(P-2, 0) each).
Every rank therefore positions its own explicit stream by skipping the draws of the images
that precede its block -- a pure host computation on the [B][C] presence matrix, which every
rank either holds or can rebuild from the image-level labels.
"""
import numpy as np

from . import _lib
from .plan import OrderStream, draws_for


def shard_range(n_images, world_size, rank):
    """Contiguous block [lo, hi) of images for `rank`; sizes differ by at most one."""
    base, extra = divmod(int(n_images), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def draws_before(present, num_iter, lo):
    """rand() draws consumed by images [0, lo) of a batch with presence matrix `present`."""
    P = np.asarray(present, dtype=np.int64).astype(bool).sum(axis=1)
    return int(sum(draws_for(int(p), num_iter) for p in P[:lo]))


def positioned_stream(present, num_iter, lo, seed=1, start_draws=0):
    """OrderStream standing where a single sequential run over the whole batch would stand
    when it reaches image `lo` (the stream started at `seed`, `start_draws` draws in)."""
    s = OrderStream(seed)
    s.skip(start_draws + draws_before(present, num_iter, lo))
    return s


def draw_order(present, num_iter, stream):
    """Class visiting order [B][num_iter][C] (-1 padded) of consecutive images on `stream`
    (weaklabel_draw_order_host; host only, no device needed)."""
    import ctypes as C
    L = _lib.load()
    p = np.ascontiguousarray(present, dtype=np.uint8)
    B, Cc = p.shape
    out = np.full((B, num_iter, Cc), -1, np.int32)
    _lib.check(L.weaklabel_draw_order_host(p.ctypes.data_as(C.POINTER(C.c_ubyte)), B, Cc, int(num_iter),
                                           stream.handle if stream is not None else None,
                                           out.ctypes.data_as(C.POINTER(C.c_int))))
    return out
