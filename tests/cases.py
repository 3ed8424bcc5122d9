"""tests/cases.py -- seeded parity cases for the EM-Adapt E-step.

Every case is reproducible from its name alone: scores and ground-truth label maps are
drawn from numpy Generators with fixed seeds, so the golden generator (which needs
/root/reference), the CPU oracle tests and the GPU parity tests all see identical bytes.
"""
import hashlib

import numpy as np

C21 = 21
DEFAULTS = dict(suppress_others=True, num_iter=5, margin_others=1e-5, bg=0.4, fg=0.2)


# ----------------------------------------------------------------------------------------
# ground-truth label maps
# ----------------------------------------------------------------------------------------
def paint_label_map(H, W, classes, rng, border255=True):
    """GT class map [H,W] int32 whose stride-2 scan (weaklabel.cpp:24-25) sees exactly
    `classes`: the first class is the fill, every other class gets a rectangle, and each
    class is finally stamped on its own even (h, w) anchor; a 1-pixel border of 255 (void)
    is added on request (it covers even coordinates too and must be ignored)."""
    lab = np.full((H, W), classes[0], np.int32)
    lo = 1 if border255 else 0
    eh = np.arange(lo + (lo & 1), H - lo, 2)
    ew = np.arange(lo + (lo & 1), W - lo, 2)
    cells = [(int(h), int(w)) for h in eh for w in ew]
    assert len(cells) >= len(classes), "image too small for this label set"
    pick = rng.choice(len(cells), size=len(classes), replace=False)
    anchors = [cells[int(i)] for i in pick]
    for k, (h0, w0) in zip(classes[1:], anchors[1:]):
        hh = int(rng.integers(1, max(2, H // 3)))
        ww = int(rng.integers(1, max(2, W // 3)))
        lab[h0:min(H - lo, h0 + hh), w0:min(W - lo, w0 + ww)] = k
    for k, (h0, w0) in zip(classes, anchors):
        lab[h0, w0] = k
    if border255:
        lab[0, :] = 255; lab[-1, :] = 255; lab[:, 0] = 255; lab[:, -1] = 255
    return lab


def voc_label_sets(B, rng, forced=None):
    """VOC-style image label sets: background + k in {0,1,2,3} foreground classes with
    P(k) = .1/.5/.3/.1.  `forced` maps image index -> explicit class list."""
    sets = []
    for b in range(B):
        k = int(rng.choice(4, p=[0.1, 0.5, 0.3, 0.1]))
        fgc = sorted(int(x) for x in rng.choice(np.arange(1, C21), size=k, replace=False))
        sets.append([0] + fgc)
    for b, s in (forced or {}).items():
        sets[b] = list(s)
    return sets


def label_maps_for_sets(sets, H, W, rng, border255=True):
    return np.stack([paint_label_map(H, W, s, rng, border255) for s in sets]).astype(np.int32)


# ----------------------------------------------------------------------------------------
# score maps
# ----------------------------------------------------------------------------------------
def bilinear_upsample(x, Ho, Wo):
    """tf.image.resize_bilinear(align_corners=False) restated in numpy (network.py:40):
    src = dst * (in/out), lerp between floor and floor+1 clamped."""
    B, H, W, Cc = x.shape
    ys = np.arange(Ho, dtype=np.float32) * np.float32(H / Ho)
    xs = np.arange(Wo, dtype=np.float32) * np.float32(W / Wo)
    y0 = np.floor(ys).astype(np.int64); x0 = np.floor(xs).astype(np.int64)
    y1 = np.minimum(y0 + 1, H - 1); x1 = np.minimum(x0 + 1, W - 1)
    fy = (ys - y0).astype(np.float32)[None, :, None, None]
    fx = (xs - x0).astype(np.float32)[None, None, :, None]
    top = x[:, y0][:, :, x0] * (1 - fx) + x[:, y0][:, :, x1] * fx
    bot = x[:, y1][:, :, x0] * (1 - fx) + x[:, y1][:, :, x1] * fx
    return (top * (1 - fy) + bot * fy).astype(np.float32)


def scores(kind, B, H, W, Cc, seed, scale=3.0):
    rng = np.random.default_rng(seed)
    shp = (B, H, W, Cc)
    if kind == "randn":
        return (rng.standard_normal(shp, dtype=np.float32) * np.float32(scale)).astype(np.float32)
    if kind == "neg":          # every score negative: exposes the FLT_MIN floor (weaklabel.cpp:71)
        return (-np.abs(rng.standard_normal(shp, dtype=np.float32)) * np.float32(scale)
                - np.float32(0.5)).astype(np.float32)
    if kind == "mixedneg":     # about half of the pixels have all-negative present scores
        return (rng.standard_normal(shp, dtype=np.float32) * np.float32(scale)
                - np.float32(1.0)).astype(np.float32)
    if kind == "int":          # integer-valued scores: heavy ties in every margin
        return rng.integers(-3, 4, size=shp).astype(np.float32)
    if kind == "const":
        return np.full(shp, np.float32(scale), np.float32)
    if kind == "dup":          # duplicated score columns: exact ties between classes
        x = rng.standard_normal(shp, dtype=np.float32) * np.float32(scale)
        x[..., 1::2] = x[..., 0:-1:2][..., : x[..., 1::2].shape[-1]]
        return x.astype(np.float32)
    if kind == "big":          # large magnitude: margin_others is absorbed, shift rounding merges
        return (rng.standard_normal(shp, dtype=np.float32) * np.float32(1e4)
                + np.float32(3e4)).astype(np.float32)
    if kind == "tiny":
        return (rng.standard_normal(shp, dtype=np.float32) * np.float32(1e-30)).astype(np.float32)
    if kind == "coarse":       # quantised to 1/8: ties in margins AND rounding-merge pressure
        return (np.round(rng.standard_normal(shp, dtype=np.float32) * 8) / 8).astype(np.float32)
    if kind == "upsampled":    # config 3: bilinear upsample of a 41x41 fc8 tensor
        small = rng.standard_normal((B, 41, 41, Cc), dtype=np.float32) * np.float32(scale)
        return bilinear_upsample(small, H, W)
    raise KeyError(kind)


def restride(x, how):
    """Return a float32 view with the same values but reference-legal non-trivial strides
    (weaklabel.cpp:50 honours stride[0] and stride[2] only)."""
    B, H, W, Cc = x.shape
    if how == "contig":
        return np.ascontiguousarray(x)
    if how == "padpix":        # padded pixel stride: stride[2] = (C+3)*4
        buf = np.full((B, H, W, Cc + 3), np.float32(777.0), np.float32)
        buf[..., :Cc] = x
        return buf[..., :Cc]
    if how == "padbatch":      # batch stride larger than H*W*C*4
        buf = np.full((B, H * W * Cc + 100), np.float32(555.0), np.float32)
        buf[:, :H * W * Cc] = x.reshape(B, -1)
        return buf[:, :H * W * Cc].reshape(B, H, W, Cc)
    if how == "offset1":       # 4-byte aligned but not 16-byte aligned base
        buf = np.zeros(B * H * W * Cc + 1, np.float32)
        v = buf[1:].reshape(B, H, W, Cc)
        v[...] = x
        return v
    raise KeyError(how)


# ----------------------------------------------------------------------------------------
# the case table
# ----------------------------------------------------------------------------------------
def _case(name, B, H, W, kind="randn", seed=0, scale=3.0, sets=None, stride="contig",
          label_fn=None, store="labels", **params):
    p = dict(DEFAULTS); p.update(params)
    return dict(name=name, B=B, H=H, W=W, C=C21, kind=kind, seed=seed, scale=scale, sets=sets,
                stride=stride, label_fn=label_fn, store=store, params=p)


def _lab_odd_only(H, W):
    """class 7 only at odd coordinates (dropped), class 9 only at (H-1,W-1) (even for odd H: kept)."""
    lab = np.zeros((1, H, W), np.int32)
    lab[0, 1::2, 1::2] = 7
    lab[0, H - 1, W - 1] = 9
    return lab


def _lab_void(H, W):
    """GT with 255, 21, 37 sprinkled on even coordinates; classes 0, 3, 15 present."""
    rng = np.random.default_rng(77)
    lab = paint_label_map(H, W, [0, 3, 15], rng, border255=True)[None].copy()
    lab[0, 4, 4] = 21; lab[0, 6, 8] = 37; lab[0, 10, 10] = 255; lab[0, 12, 2] = 2 ** 20
    return lab


CASES = [
    # (1) config 1: one 41x41x21 map, label set {0,k}: no randomness consumed
    _case("c1_41_p2", 1, 41, 41, seed=101, sets=[[0, 15]], store="full"),
    # (2) P = 1
    _case("p1_bg_only", 1, 41, 41, seed=102, sets=[[0]]),
    _case("p1_fg_only", 2, 41, 41, seed=103, sets=[[12], [5]]),
    _case("p2_nobg", 1, 41, 41, seed=104, sets=[[3, 17]]),
    # (3) rand() carry-over and cumulative permutation across consecutive images
    _case("p5_seq3", 3, 41, 41, seed=105, sets=[[0, 2, 7, 11, 19], [0, 1, 5, 9, 20], [4, 6, 8, 10, 12]], store="full"),
    _case("p3_seq4", 4, 33, 47, seed=106, sets=[[0, 1, 2], [0, 8, 9], [0, 3, 20], [1, 2, 3]]),
    # (4) all 21 classes
    _case("p21_41", 1, 41, 41, seed=107, sets=[list(range(21))]),
    _case("p20_nobg_24", 1, 24, 30, seed=108, sets=[list(range(1, 21))]),
    # (5) stride-2 label scan corner cases
    _case("odd_only", 1, 41, 41, seed=109, label_fn="odd_only"),
    # (6) void labels
    _case("void_labels", 1, 41, 41, seed=110, label_fn="void"),
    # (7) FLT_MIN floor
    _case("allneg_p1", 1, 41, 41, kind="neg", seed=111, sets=[[0]], store="full"),
    _case("allneg_p3", 2, 41, 41, kind="neg", seed=112, sets=[[0, 4, 9], [0, 13]]),
    _case("mixedneg_p2", 2, 41, 41, kind="mixedneg", seed=113, sets=[[0, 6], [7]]),
    # (8) exact ties
    _case("ties_int", 2, 41, 41, kind="int", seed=114, sets=[[0, 5, 6], [0, 1]], store="full"),
    _case("ties_const", 1, 21, 21, kind="const", seed=115, scale=1.5, sets=[[0, 2, 9]], store="full"),
    _case("ties_dup", 2, 41, 41, kind="dup", seed=116, sets=[[0, 1, 2, 3], [2, 3]]),
    _case("ties_coarse", 2, 41, 41, kind="coarse", seed=117, sets=[[0, 10], [0, 3, 4, 18]]),
    # (9) strides
    _case("stride_padpix", 2, 41, 41, seed=118, sets=[[0, 14], [0, 2, 3]], stride="padpix"),
    _case("stride_padbatch", 3, 41, 41, seed=119, sets=[[0, 14], [0, 2, 3], [0]], stride="padbatch"),
    _case("stride_offset1", 2, 41, 41, seed=120, sets=[[0, 11], [0, 4, 7]], stride="offset1"),
    # (10) parameters
    _case("nosuppress_p3", 2, 41, 41, seed=121, sets=[[0, 5, 9], [3]], suppress_others=False, store="full"),
    _case("iter0", 1, 41, 41, seed=122, sets=[[0, 5, 9]], num_iter=0, store="full"),
    _case("iter1", 1, 41, 41, seed=123, sets=[[0, 5, 9, 13]], num_iter=1),
    _case("iter9", 1, 41, 41, seed=124, sets=[[0, 5, 9, 13]], num_iter=9),
    _case("rank_first", 1, 41, 41, seed=125, sets=[[0, 8]], bg=0.0, fg=0.0),
    _case("rank_last", 1, 41, 41, seed=126, sets=[[0, 8]], bg=0.9999, fg=0.9999),
    _case("defaults_py", 1, 41, 41, seed=127, sets=[[0, 1, 2]], bg=0.5, fg=0.25),
    _case("margin_big", 1, 41, 41, seed=128, sets=[[0, 6]], margin_others=0.5),
    # a negative margin lifts suppressed absent classes ABOVE the smallest present one
    _case("margin_neg", 2, 41, 41, seed=138, sets=[[0, 6, 11], [4]], margin_others=-0.75, store="full"),
    # (12) magnitudes
    _case("mag_big", 2, 41, 41, kind="big", seed=129, sets=[[0, 9], [0, 3, 5]]),
    _case("mag_big_p1", 1, 41, 41, kind="big", seed=130, sets=[[7]], store="full"),
    _case("mag_tiny", 1, 41, 41, kind="tiny", seed=131, sets=[[0, 9, 10]]),
    _case("mag_scale5", 2, 41, 41, seed=132, scale=5.0, sets=[[0, 12, 15], [0, 20]]),
    # ragged / tiny shapes
    _case("shape_1x1", 1, 1, 1, seed=133, sets=[[0]], label_fn="zeros", store="full"),
    _case("shape_2x3", 1, 2, 3, seed=134, sets=[[0]], label_fn="zeros", store="full"),
    _case("shape_7x5", 2, 7, 5, seed=135, label_fn="zeros", store="full"),
    _case("shape_64x64", 1, 64, 64, seed=136, sets=[[0, 2, 19]]),
    _case("shape_100x37", 2, 100, 37, seed=137, sets=[[0, 1], [0, 5, 6, 7]]),
    # (11) the BASELINE.json configurations
    _case("c2_41_b30", 30, 41, 41, seed=201, sets="voc_c2", store="labels"),
    _case("c3_321_b30", 30, 321, 321, kind="upsampled", seed=202, sets="voc", store="sha"),
    _case("c3_321_randn", 3, 321, 321, seed=203, sets="voc", store="sha"),
    _case("c4_513_b4", 4, 513, 513, seed=204, sets="voc", store="sha"),
    _case("c4_513_p21", 1, 513, 513, seed=205, sets=[list(range(21))], store="sha"),
    _case("c4_513_scale5", 2, 513, 513, seed=206, scale=5.0, sets=[[0, 1, 2, 3, 4], [0, 15]], store="sha"),
]
CASE_BY_NAME = {c["name"]: c for c in CASES}
# cases the CPU suite runs end to end through the C restatement (the big ones are slow-ish
# but still seconds); GPU parity runs all of them.
SMALL = [c["name"] for c in CASES if c["H"] * c["W"] * c["B"] <= 41 * 41 * 30]


def build_case(case):
    """-> (fmap float32 [B,H,W,C] with the case's strides, label int32 [B,H,W], params)"""
    if isinstance(case, str):
        case = CASE_BY_NAME[case]
    B, H, W, Cc = case["B"], case["H"], case["W"], case["C"]
    x = scores(case["kind"], B, H, W, Cc, case["seed"], case["scale"])
    lrng = np.random.default_rng(4321 + case["seed"])
    fn = case["label_fn"]
    if fn == "odd_only":
        lab = _lab_odd_only(H, W)
    elif fn == "void":
        lab = _lab_void(H, W)
    elif fn == "zeros":
        lab = np.zeros((B, H, W), np.int32)
    else:
        sets = case["sets"]
        if sets == "voc":
            sets = voc_label_sets(B, lrng)
        elif sets == "voc_c2":       # config 2: image 0 background only, image 1 all classes
            sets = voc_label_sets(B, lrng, forced={0: [0], 1: list(range(21))})
        lab = label_maps_for_sets(sets, H, W, lrng, border255=min(H, W) >= 8)
    return restride(x, case["stride"]), np.ascontiguousarray(lab, np.int32), dict(case["params"])


def sha1(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()
