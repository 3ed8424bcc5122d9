"""Parity tests proper (B200): libweaklabel_cuda.so, called through its C-ABI (ctypes), against
the oracle on the same seeded inputs and against the committed golden fixtures (outputs of the
reference's own weaklabel.cpp).  Bit-exact for labels, maps, biases and the two means; 1e-5
relative for the softmax-CE loss and gradient (the tolerance BASELINE.json's north_star states).
Nothing here reads /root/reference."""
import ctypes

import numpy as np
import pytest

import cases
import em_adapt_tensorflow_b200 as wl
from oracle.oracle import Oracle

import os

pytestmark = pytest.mark.gpu
ALL = [c["name"] for c in cases.CASES]


def _torch():
    import torch
    return torch


def _oracle_run(f, l, p):
    orc = Oracle(1)
    fo = np.array(f, dtype=np.float32, order="C", copy=True)
    probes = orc.e_step_inplace(fo, l, probes=True, **p)
    return fo, orc.argmax_labels(fo), probes


def _fast_applies(p):
    return bool(p["suppress_others"]) and p["margin_others"] >= 0


def _device_run(f, l, p, tune=None, want_map=True, path="auto"):
    torch = _torch()
    B, H, W, C = f.shape
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    plan = wl.EStepPlan(B, H, W, C)
    if tune:
        plan.tune(*tune)
    plan.set_path(path)
    plan.prepare(wl.label_set(l, C), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                 rng=wl.OrderStream(1))
    lab = plan.labels(t)
    torch.cuda.synchronize()
    assert t.cpu().numpy().tobytes() == np.ascontiguousarray(f).tobytes(), "labels() must not modify the map"
    probes = plan.probes()
    if path == "auto" and tune is None:
        assert plan.last_path == ("fast" if _fast_applies(p) else "generic")
    elif path != "auto":
        assert plan.last_path == path
    fm = None
    if want_map:
        plan.finalize_map(t)
        fm = t.cpu().numpy()
    plan.close()
    return fm, lab.cpu().numpy(), probes


@pytest.mark.parametrize("path", ["auto", "generic"])
@pytest.mark.parametrize("name", ALL)
def test_device_api_bit_exact(name, path, golden):
    """path auto = the on-chip fast kernel wherever it applies; generic = estep_kernel.cu."""
    meta, arrays = golden
    f, l, p = cases.build_case(name)
    want_map, want_lab, (order_o, bias_o, m0_o, m1_o) = _oracle_run(f, l, p)
    got_map, got_lab, (order_g, bias_g, m0_g, m1_g) = _device_run(f, l, p, path=path)
    assert np.array_equal(order_g, order_o)
    assert bias_g.tobytes() == bias_o.tobytes()
    assert m0_g.tobytes() == m0_o.tobytes() and m1_g.tobytes() == m1_o.tobytes()
    assert np.array_equal(got_lab, want_lab)
    assert got_map.tobytes() == want_map.tobytes()
    # and against the reference's own outputs
    assert cases.sha1(got_lab) == meta[name]["labels_sha1"]
    assert cases.sha1(got_map) == meta[name]["map_sha1"]


@pytest.mark.parametrize("name", [n for n in ALL if _fast_applies(cases.CASE_BY_NAME[n]["params"])])
def test_labels_only_mode_gives_the_same_labels(name, golden):
    """weaklabel_plan_set_outputs(plan, 0): the two order-exact sums are formed only for images in which a
    pixel's argmax can depend on the rounding of the re-centring shift (heavy ties, huge magnitudes ...);
    the labels must be the reference's in every case, and images whose sums were formed carry exact ones."""
    torch = _torch()
    meta, _ = golden
    f, l, p = cases.build_case(name)
    if f.strides[2] != f.shape[3] * 4:
        pytest.skip("padded pixel stride: generic kernel")
    want_map, want_lab, (order_o, bias_o, m0_o, m1_o) = _oracle_run(f, l, p)
    B, H, W, C = f.shape
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    plan = wl.EStepPlan(B, H, W, C)
    plan.set_path("fast").set_outputs(sums=False)
    plan.prepare(wl.label_set(l, C), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"], rng=wl.OrderStream(1))
    lab = plan.labels(t)
    torch.cuda.synchronize()
    assert np.array_equal(lab.cpu().numpy(), want_lab)
    assert cases.sha1(lab.cpu().numpy()) == meta[name]["labels_sha1"]
    _, bias_g, m0_g, m1_g = plan.probes()
    assert bias_g.tobytes() == bias_o.tobytes()
    done = ~np.isnan(m1_g)
    assert m0_g[done].tobytes() == m0_o[done].tobytes() and m1_g[done].tobytes() == m1_o[done].tobytes()
    with pytest.raises(wl.WeakLabelError):
        plan.finalize_map(t)                               # needs the sums of every image
    plan.set_outputs(sums=True)
    assert torch.equal(plan.labels(t), lab)
    plan.finalize_map(t)
    assert t.cpu().numpy().tobytes() == want_map.tobytes()
    plan.close()


@pytest.mark.parametrize("name", [n for n in ALL if cases.CASE_BY_NAME[n]["H"] <= 100])
def test_legacy_e_step_bit_exact(name, golden):
    """The drop-in symbol on HOST buffers with the case's own (possibly padded) strides and the
    process-global rand() stream, exactly as estep.py:29-31 calls it."""
    meta, _ = golden
    f, l, p = cases.build_case(name)
    ctypes.CDLL("libc.so.6").srand(1)
    out = wl.estep_c(f, l, p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"])
    assert out is f
    assert cases.sha1(np.ascontiguousarray(f)) == meta[name]["map_sha1"]
    assert cases.sha1(np.argmax(f, axis=3).astype(np.uint8)) == meta[name]["labels_sha1"]


def test_legacy_padding_untouched():
    f, l, p = cases.build_case("stride_padpix")
    base = f.base
    wl.estep_c(f, l, True, 5, 1e-5, 0.4, 0.2)
    assert (base[..., 21:] == np.float32(777.0)).all()
    f, l, p = cases.build_case("stride_padbatch")
    base = f.base if f.base is not None else f
    while getattr(base, "base", None) is not None:
        base = base.base
    wl.estep_c(f, l, True, 5, 1e-5, 0.4, 0.2)
    assert (np.asarray(base).reshape(3, -1)[:, 41 * 41 * 21:] == np.float32(555.0)).all()


def test_estep_wrapper_matches_reference_module_contract():
    """estep.py:18-27: returns a new float32 array, inputs untouched, label may be float."""
    f, l, p = cases.build_case("c1_41_p2")
    f0 = f.copy()
    ctypes.CDLL("libc.so.6").srand(1)
    out = wl.estep(f, l.astype(np.float32), True, 5, 1e-5, 0.4, 0.2, use_c=True)
    assert out.dtype == np.float32 and out is not f and np.array_equal(f, f0)
    want, _, _ = _oracle_run(f, l, p)
    assert out.tobytes() == want.tobytes()


TUNES = [(1, 64, 0), (2, 64, 0), (4, 128, 0), (8, 32, 2), (2, 0, 0), (16, 16, 0)]


@pytest.mark.parametrize("tune", TUNES)
@pytest.mark.parametrize("name", ["c1_41_p2", "p5_seq3", "nosuppress_p3", "p21_41", "allneg_p3", "ties_int",
                                  "shape_7x5", "mag_big"])
def test_forced_geometries(name, tune):
    """Generic kernel: cluster sizes 1..16, multi-tile and multi-round geometries give the same bits."""
    f, l, p = cases.build_case(name)
    want_map, want_lab, (order_o, bias_o, m0_o, m1_o) = _oracle_run(f, l, p)
    got_map, got_lab, (order_g, bias_g, m0_g, m1_g) = _device_run(f, l, p, tune=tune, path="generic")
    assert bias_g.tobytes() == bias_o.tobytes()
    assert m0_g.tobytes() == m0_o.tobytes() and m1_g.tobytes() == m1_o.tobytes()
    assert np.array_equal(got_lab, want_lab)
    assert got_map.tobytes() == want_map.tobytes()


FAST_TUNES = [(1, 0, 0), (2, 0, 0), (4, 0, 0), (8, 0, 1), (9, 0, 0), (3, 0, 2), (2, 1200, 0), (1, 4100, 0), (2, 5000, 2)]


@pytest.mark.parametrize("tune", FAST_TUNES)
@pytest.mark.parametrize("name", ["c1_41_p2", "p5_seq3", "p21_41", "allneg_p3", "ties_int", "ties_const", "shape_7x5",
                                  "shape_1x1", "mag_big", "mag_tiny", "rank_last", "c2_41_b30", "shape_100x37"])
def test_fast_path_forced_geometries(name, tune):
    """On-chip kernel (estep_onchip.cu): cluster sizes 1..9 (DSMEM exchange), tiles larger than the image
    (idle CTAs), the per-thread quad counts (small tiles select the small builds)."""
    f, l, p = cases.build_case(name)
    N = f.shape[1] * f.shape[2]
    G, TS, _ = tune
    if TS and G * TS < N:
        pytest.skip("tile tuning does not hold the image")
    want_map, want_lab, (order_o, bias_o, m0_o, m1_o) = _oracle_run(f, l, p)
    got_map, got_lab, (order_g, bias_g, m0_g, m1_g) = _device_run(f, l, p, tune=tune, path="fast")
    assert bias_g.tobytes() == bias_o.tobytes()
    assert m0_g.tobytes() == m0_o.tobytes() and m1_g.tobytes() == m1_o.tobytes()
    assert np.array_equal(got_lab, want_lab)
    assert got_map.tobytes() == want_map.tobytes()


@pytest.mark.parametrize("name,tune", [("c3_321_b30", (0, 0, 2)), ("c3_321_b30", (0, 0, 7)), ("c4_513_b4", (9, 0, 1)),
                                       ("c4_513_b4", (9, 0, 2)), ("c4_513_scale5", (9, 0, 1)), ("c2_41_b30", (9, 0, 3)),
                                       ("c4_513_p21", (9, 0, 1)), ("p21_41", (3, 0, 1)), ("c3_321_b30", (5, 0, 3)),
                                       ("c3_321_b30", (7, 0, 4)), ("c2_41_b30", (6, 0, 2))])
def test_onchip_kernel_cycles_images(name, tune):
    """The on-chip kernel with FEWER resident clusters than images: every cluster works through a queue of
    images while its producer warps stream the next one (staging-ring reuse across images, both plane sets
    of the scratch, command / ready hand-shake), incl. the non-portable cluster sizes 3, 5, 6, 7, 9."""
    f, l, p = cases.build_case(name)
    want_map, want_lab, (order_o, bias_o, m0_o, m1_o) = _oracle_run(f, l, p)
    for _ in range(2):
        got_map, got_lab, (order_g, bias_g, m0_g, m1_g) = _device_run(f, l, p, tune=tune, path="fast")
        assert bias_g.tobytes() == bias_o.tobytes()
        assert m0_g.tobytes() == m0_o.tobytes() and m1_g.tobytes() == m1_o.tobytes()
        assert np.array_equal(got_lab, want_lab)
        assert got_map.tobytes() == want_map.tobytes()


def test_fast_path_refuses_what_it_does_not_cover():
    torch = _torch()
    f, l, p = cases.build_case("nosuppress_p3")
    plan = wl.EStepPlan(*f.shape)
    plan.set_path("fast")
    plan.prepare(wl.label_set(l, 21), suppress_others=False, rng=wl.OrderStream(1))
    with pytest.raises(wl.WeakLabelError):
        plan.labels(torch.from_numpy(np.ascontiguousarray(f)).cuda())
    plan.close()


def test_one_call_device_entry_and_repeatability():
    """e_step_labels_device (cached plan, host GT map, device scores) twice on explicit streams
    from the same seed gives identical labels; the score map is left untouched."""
    torch = _torch()
    import ctypes as C
    f, l, p = cases.build_case("p3_seq4")
    B, H, W, Cc = f.shape
    L = wl.load()
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    shape = (C.c_int * 4)(B, H, W, Cc)
    stride = (C.c_int * 4)(*[s * 4 for s in t.stride()])
    outs = []
    for _ in range(2):
        lab = torch.empty((B, H, W), dtype=torch.uint8, device="cuda")
        rng = wl.OrderStream(1)
        rc = L.e_step_labels_device(t.data_ptr(), shape, stride, l.ctypes.data_as(C.POINTER(C.c_int)), True, 5,
                                    1e-5, 0.4, 0.2, rng.handle, lab.data_ptr(), None, None)
        assert rc == 0, L.weaklabel_last_error()
        torch.cuda.synchronize()
        outs.append(lab.cpu().numpy())
    _, want_lab, _ = _oracle_run(f, l, p)
    assert np.array_equal(outs[0], want_lab) and np.array_equal(outs[1], want_lab)
    assert t.cpu().numpy().tobytes() == np.ascontiguousarray(f).tobytes()


def test_sharded_batch_equals_whole_batch():
    """Two 'ranks' on one device: each runs its image block with a positioned order stream;
    concatenated labels equal the single-stream run (no collective on the path)."""
    torch = _torch()
    f, l, p = cases.build_case("c2_41_b30")
    _, want_lab, _ = _oracle_run(f, l, p)
    present = wl.label_set(l, 21)
    parts = []
    for rank in range(2):
        lo, hi = wl.shard_range(30, 2, rank)
        t = torch.from_numpy(np.ascontiguousarray(f[lo:hi])).cuda()
        plan = wl.EStepPlan(hi - lo, 41, 41, 21)
        plan.prepare(present[lo:hi], rng=wl.positioned_stream(present, 5, lo))
        parts.append(plan.labels(t).cpu().numpy())
        plan.close()
    assert np.array_equal(np.concatenate(parts), want_lab)


def test_errors_are_reported_not_swallowed():
    torch = _torch()
    plan = wl.EStepPlan(1, 8, 8, 21)
    with pytest.raises(wl.WeakLabelError):                 # empty label set (reference: crash)
        plan.prepare(np.zeros((1, 21), np.uint8))
    with pytest.raises(wl.WeakLabelError):                 # rank outside [0, N)
        plan.prepare(np.eye(1, 21, dtype=np.uint8), bg=1.5)
    t = torch.zeros((1, 8, 8, 21), device="cuda")
    with pytest.raises(wl.WeakLabelError):                 # not prepared
        plan.labels(t)
    plan.close()


OC_FLAGS = {"rankcount_off": 1, "resident_off": 2, "fastlabel_off": 4, "loose_zeros": 8, "digest_off": 16,
            "classic_xchg": 32, "ragged_copies": 64, "all_off": 127}


@pytest.mark.parametrize("flag", sorted(OC_FLAGS))
@pytest.mark.parametrize("name,tune", [("p5_seq3", (0, 0, 0)), ("c2_41_b30", (2, 0, 3)), ("ties_int", (4, 0, 0)),
                                       ("mag_tiny", (2, 0, 0)), ("allneg_p3", (0, 0, 0)), ("c3_321_randn", (0, 0, 0)),
                                       ("c4_513_b4", (9, 0, 1))])
def test_fast_kernel_switches(name, tune, flag, monkeypatch):
    """Every optimisation of estep_onchip.cu has a switch (WEAKLABEL_FLAGS, estep_kernel.cuh OC_F_*) that
    routes around it: rank-counting selection, planes resident in tensor memory, float label path, tight
    zero-count bound, pushed sum digests, st.async exchanges.  All combinations give the same bits."""
    monkeypatch.setenv("WEAKLABEL_FLAGS", str(OC_FLAGS[flag]))
    f, l, p = cases.build_case(name)
    want_map, want_lab, (order_o, bias_o, m0_o, m1_o) = _oracle_run(f, l, p)
    got_map, got_lab, (order_g, bias_g, m0_g, m1_g) = _device_run(f, l, p, tune=tune if any(tune) else None, path="fast")
    assert bias_g.tobytes() == bias_o.tobytes()
    assert m0_g.tobytes() == m0_o.tobytes() and m1_g.tobytes() == m1_o.tobytes()
    assert np.array_equal(got_lab, want_lab)
    assert got_map.tobytes() == want_map.tobytes()


def test_fast_kernel_soak():
    """Hundreds of back-to-back launches of the fast kernel over varying resident-cluster counts, cluster
    sizes, two streams and alternating shapes, labels checked every time (compute-sanitizer is closed on
    this pool, so hazards between launches / images / mbarrier phases are hunted by repetition)."""
    torch = _torch()
    jobs = []
    for name, tunes in [("c2_41_b30", [(1, 0, 1), (1, 0, 7), (2, 0, 2), (3, 0, 15), (9, 0, 1), (4, 0, 0)]),
                        ("p5_seq3", [(1, 0, 0), (2, 0, 1), (5, 0, 0)]),
                        ("c3_321_randn", [(0, 0, 0), (4, 0, 1), (7, 0, 0)]),
                        ("shape_100x37", [(1, 0, 0), (2, 0, 0), (6, 0, 1)])]:
        f, l, p = cases.build_case(name)
        _, want_lab, _ = _oracle_run(f, l, p)
        t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
        want = torch.from_numpy(want_lab).cuda()
        for tune in tunes:
            plan = wl.EStepPlan(*f.shape)
            plan.tune(*tune)
            plan.set_path("fast")
            plan.prepare(wl.label_set(l, f.shape[3]), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"],
                         p["fg"], rng=wl.OrderStream(1))
            jobs.append((plan, t, want, torch.empty_like(want)))
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    n = 0
    for rep in range(36):
        for j, (plan, t, want, out) in enumerate(jobs):
            with torch.cuda.stream(streams[(rep + j) & 1]):
                plan.labels(t, out=out)
            n += 1
        if rep % 6 == 5:
            torch.cuda.synchronize()
            for plan, t, want, out in jobs:
                assert torch.equal(out, want)
    torch.cuda.synchronize()
    for plan, t, want, out in jobs:
        assert torch.equal(out, want)
        plan.close()
    assert n >= 500


# ---------------------------------------------------------------------------------------------
# size-independent properties at BASELINE.json's full image size (513x513, several images)
# ---------------------------------------------------------------------------------------------
def test_full_size_properties():
    torch = _torch()
    B, H, W, C = 6, 513, 513, 21
    rng = np.random.default_rng(8)
    sets = cases.voc_label_sets(B, rng, forced={0: [0], 1: [0, 3, 7, 11, 15]})
    present = np.zeros((B, C), np.uint8)
    for b, s in enumerate(sets):
        present[b, s] = 1
    g = torch.Generator(device="cuda")
    g.manual_seed(77)
    t = torch.empty((B, H, W, C), device="cuda").normal_(0, 3, generator=g)
    plan = wl.EStepPlan(B, H, W, C)
    plan.prepare(present, rng=wl.OrderStream(1))
    lab = plan.labels(t)
    lab2 = plan.labels(t)                                  # idempotent on an unmodified map
    assert torch.equal(lab, lab2)
    order, bias, m0, m1 = plan.probes()
    assert (bias >= 0).all()                               # margins are >= 0 (weaklabel.cpp:86)
    N = H * W
    labn = lab.cpu().numpy().reshape(B, N)
    for b, s in enumerate(sets):
        assert set(np.unique(labn[b])) <= set(s)           # absent classes are suppressed
        hist = np.bincount(labn[b], minlength=C)
        last = order[b, -1][order[b, -1] >= 0]
        c_last = int(last[-1])                             # the class visited last holds its quota
        nth = int(np.float32(0.4 if c_last == 0 else 0.2) * np.float32(N))
        assert hist[c_last] >= nth + 1 - 8                 # (ties at the pivot can go either way by index)
    # the finalized map's argmax is the label map, and re-centring keeps the mean max (weaklabel.cpp:102-112)
    t2 = t.clone()
    plan.finalize_map(t2)
    assert torch.equal(t2.argmax(dim=3).to(torch.uint8), lab)
    shift = torch.from_numpy(m0.astype(np.float64) - m1).cuda().view(B, 1, 1)
    floored = (t2.amax(dim=3).double() - shift).clamp_min(float(np.finfo(np.float32).tiny))
    assert np.allclose(floored.mean(dim=(1, 2)).cpu().numpy(), m1, rtol=0, atol=1e-4)
    plan.close()


# ---------------------------------------------------------------------------------------------
# fused softmax cross-entropy (deeplab.py:182, ignore semantics of network.py:77-89)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,ignore_frac", [(1, 0.0), (41 * 41 * 3, 0.0), (41 * 41 * 30, 0.3), (1000, 1.0),
                                           (321 * 321 * 2 + 5, 0.1)])
def test_softmax_ce_matches_oracle(n, ignore_frac, oracle):
    torch = _torch()
    rng = np.random.default_rng(n)
    z = (rng.standard_normal((n, 21)) * 4).astype(np.float32)
    y = rng.integers(0, 21, n).astype(np.uint8)
    y[rng.random(n) < ignore_frac] = 255
    want_loss, kept, want_grad = oracle.softmax_ce(z, y)
    loss, valid, grad = wl.softmax_ce(torch.from_numpy(z).cuda(), torch.from_numpy(y).cuda())
    assert int(valid.item()) == kept
    tol = 1e-5                                             # north_star: within 1e-5 relative
    assert abs(float(loss.item()) - want_loss) <= tol * max(abs(want_loss), 1e-30)
    g = grad.cpu().numpy().astype(np.float64)
    assert np.abs(g - want_grad).max() <= tol * max(np.abs(want_grad).max(), 1e-30)
    assert (g[y == 255] == 0).all()


def test_softmax_ce_forward_only_and_torch_reference(oracle):
    torch = _torch()
    f, l, p = cases.build_case("c3_321_randn")
    z = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    plan = wl.EStepPlan(*f.shape)
    plan.prepare(wl.label_set(l, 21), rng=wl.OrderStream(1))
    lab = plan.labels(z)
    loss, valid, grad = wl.softmax_ce(z, lab)
    loss_f, _, none = wl.softmax_ce(z, lab, want_grad=False)
    assert none is None and float(loss_f.item()) == float(loss.item())
    zz = z.reshape(-1, 21).double().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(zz, lab.reshape(-1).long(), ignore_index=255)
    ref.backward()
    assert abs(float(loss.item()) - float(ref.item())) <= 1e-5 * float(ref.item())
    assert (grad.reshape(-1, 21).double() - zz.grad).abs().max().item() <= 1e-5 * zz.grad.abs().max().item()
    plan.close()
