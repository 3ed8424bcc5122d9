#!/usr/bin/env python
"""bench.py -- E-step throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[3], the configuration the metric is quoted on): 256 synthetic
513x513x21 float32 score maps PER GPU (weak scaling; images shard across ranks with no data-path
collective), scores N(0, 3^2) drawn on the device from torch.Generator(1234 + global image
index), VOC-style image label sets (background + 0..3 foreground classes, P = .1/.5/.3/.1) from
numpy default_rng(4321), scalars suppress=True num_iter=5 margin=1e-5 bg=0.4 fg=0.2
(deeplab.py:181).  A step = one E-step over the rank's 256 resident score maps -> uint8 labels.

One JSON line on rank 0.  `value` = pixels of all ranks / max-over-ranks device time (CUDA
events on the launch stream, barrier + synchronize on both sides).  `roofline` = algorithmic
bytes (85 B/pixel: 84 B scores read once + 1 B label) / the E-step kernel's mean launch time,
against MEASURED_PEAKS.json's hbm_gbs.  `e2e` = the same pixels through the reference-facing
legacy C-ABI `e_step` on HOST (pinned) buffers, H2D + kernels + D2H inside the timed region
(`e2e.pageable` = the same from pageable numpy memory, what estep.py:21-22 hands over).
`cpu_baseline` = the reference's own weaklabel.cpp (oracle/_ref, compiled unmodified) on the
host cores over a bounded sample (+ `variants`: the shipped -O0 binary, one core, as called from
Python with the copies of estep.py:21-27).  `legs` = the other BASELINE.json configurations and
kernels, each with its own algorithmic-byte roofline: config 2 (30x41^2), config 3 (30x321^2 +
loss), the loss alone and the in-place map on the 513^2 batch, worst-case label sets.
`--impl reference` times the CPU implementation as its own arm.  Only the cpu_baseline /
reference legs and the parity gates touch oracle/.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

H = W = 513
C = 21
IMAGES_PER_GPU = 256
PARAMS = dict(suppress_others=True, num_iter=5, margin_others=1e-5, bg=0.4, fg=0.2)
PARGS = [PARAMS[k] for k in ("suppress_others", "num_iter", "margin_others", "bg", "fg")]
ALG_BYTES_PER_PIXEL = 85          # 21*4 read + 1 label byte
CE_BYTES_PER_PIXEL = 169          # 84 read + 84 gradient written + 1 label byte
MAP_BYTES_PER_PIXEL = 168         # in-place map: 84 read + 84 written
METRIC = "E-step Gpixel/s (21-class, 513^2, B200) and % of HBM roofline"
REAL_STDOUT = None


# ------------------------------------------------------------------------------------------
# synthetic workload
# ------------------------------------------------------------------------------------------
def label_sets(n_images, seed=4321):
    import cases
    return cases.voc_label_sets(n_images, np.random.default_rng(seed))


def presence_matrix(sets, classes=C):
    p = np.zeros((len(sets), classes), np.uint8)
    for b, s in enumerate(sets):
        p[b, s] = 1
    return p


def device_scores(lo, hi, device, h=H, w=W):
    import torch
    out = torch.empty((hi - lo, h, w, C), dtype=torch.float32, device=device)
    g = torch.Generator(device=device)
    for i in range(lo, hi):
        g.manual_seed(1234 + i)
        out[i - lo].normal_(0.0, 3.0, generator=g)
    return out


def host_scores(index):
    return (np.random.default_rng(1234 + index).standard_normal((1, H, W, C), dtype=np.float32)
            * np.float32(3.0))


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's weaklabel.cpp on the host cores (oracle/_ref), one image per call
# ------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """Runs in a fresh process: generate `idx` images, then time e_step on each."""
    which, idxs, sets, barrier_t, as_python = args
    import cases
    from oracle.oracle import Oracle, Reference
    if which in ("ref", "shipped"):
        ref = Reference(which)
    else:
        orc = Oracle(1)
    data = []
    for i in idxs:
        lab = cases.paint_label_map(H, W, sets[i], np.random.default_rng(99 + i))[None].astype(np.int32)
        data.append((host_scores(i), np.ascontiguousarray(lab)))
    while time.time() < barrier_t:          # crude start barrier shared by all workers
        time.sleep(0.001)
    t0 = time.time()
    for f, l in data:
        if as_python:                       # estep.py:18-27: astype copies in, in-place call, astype copy out
            f2 = f.astype(np.float32)
            l2 = l.astype(np.int32)
            ref.e_step_image_inplace(f2, l2, **PARAMS)
            f2 = f2.astype(np.float32)
        elif which in ("ref", "shipped"):
            ref.e_step_image_inplace(f, l, **PARAMS)
        else:
            orc.e_step_inplace(f, l, **PARAMS)
    return t0, time.time(), len(data)


def cpu_reference_pass(indices, sets, cores, which=None, as_python=False):
    """Wall time of the reference CPU E-step over the images `indices` on `cores` worker processes."""
    import multiprocessing as mp
    from oracle.oracle import Reference, build
    build()
    if which is None:
        which = "ref" if Reference.available("ref") else "port"
    per = [list(indices[k::cores]) for k in range(cores)]
    per = [p for p in per if p]
    gen_s = 0.9 * max(len(p) for p in per) + 3.0          # generation ~0.6 s/image/core
    start_at = time.time() + gen_s
    ctx = mp.get_context("spawn")
    with ctx.Pool(len(per)) as pool:
        res = pool.map(_cpu_worker, [(which, p, sets, start_at, as_python) for p in per])
    t0 = min(r[0] for r in res)
    t1 = max(r[1] for r in res)
    return dict(kind="reference" if which in ("ref", "shipped") else "port", which=which, wall_s=t1 - t0,
                images=sum(r[2] for r in res), cores=len(per))


def strided_sample(n, total=IMAGES_PER_GPU):
    """n image indices striding through the 256 label sets (same P mix as the GPU arm)."""
    return [int(round(k * total / n)) % total for k in range(n)]


def cpu_variants(sets, cores):
    """the shipped -O0 binary beside the -O2 rebuild, one core, and the as-called-from-Python
    figure (three copies of estep.py:21-27 around the call); small samples, a few seconds each."""
    from oracle.oracle import Reference
    out = {}
    idx1 = strided_sample(4)
    for name, which, ncores, idx, aspy in (("O2_rebuild_1core", "ref", 1, idx1, False),
                                           ("shipped_O0_1core", "shipped", 1, idx1, False),
                                           ("shipped_O0_all_cores", "shipped", cores, strided_sample(2 * cores), False),
                                           ("O2_rebuild_as_called_from_python_1core", "ref", 1, idx1, True)):
        if not Reference.available(which):
            out[name] = None
            continue
        info = cpu_reference_pass(idx, sets, ncores, which=which, as_python=aspy)
        out[name] = dict(value=info["images"] * H * W / info["wall_s"] / 1e9, unit="Gpixel/s", cores=info["cores"],
                         images=info["images"])
    return out


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksEventReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.01)

    def result(self):
        self.stop_flag = True
        self.join(timeout=1.0)
        return dict(sm_mhz=statistics.median(self.samples) if self.samples else None,
                    sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons), samples=len(self.samples))


# ------------------------------------------------------------------------------------------
def measured_peak():
    try:
        mp_ = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(mp_["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, STREAM-style copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel, images):
    """DRAM bytes of one launch from the committed ncu capture -- only when this run launches the captured
    kernel over the captured number of images (else null: the figure would not describe this run)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if t.get("kernel_name") == kernel and int(t.get("captured_images", 0)) == images:
            return t["estep_bytes_per_launch"], t.get("source")
    except Exception:
        pass
    return None, None


def run_reference_arm(args, rank):
    if rank != 0:
        return 0
    cores = len(os.sched_getaffinity(0))
    n_img = max(cores, 2 * cores if cores <= 16 else cores)
    sets = label_sets(IMAGES_PER_GPU)
    idx = strided_sample(n_img)
    times = []
    info = None
    for step in range(args.warmup + args.steps):
        info = cpu_reference_pass(idx, sets, cores)
        if step >= args.warmup:
            times.append(info["wall_s"])
    pixels = n_img * H * W
    t = sum(times) / len(times)
    val = pixels / t / 1e9
    sample = "%d of the 256 513x513x21 images per step, striding through the label sets (one image per e_step call, %d worker processes)" % (
        n_img, info["cores"])
    line = dict(impl="reference", metric=METRIC, value=val, unit="Gpixel/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=t * 1e3, higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload="256x513x513x21 E-step per GPU (bounded CPU sample per step)",
                            images_per_step=n_img, **{k: (v if not isinstance(v, bool) else int(v))
                                                      for k, v in PARAMS.items()}),
                cpu_baseline=dict(value=val, unit="Gpixel/s", cores=info["cores"], kind=info["kind"],
                                  sample=sample),
                e2e=dict(value=val, unit="Gpixel/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0)
    print(json.dumps(line), file=REAL_STDOUT or sys.stdout, flush=True)
    return 0


# ------------------------------------------------------------------------------------------
# the other configurations and kernels (rank 0, N = 1): each against its own algorithmic bytes
# ------------------------------------------------------------------------------------------
def timed(fn, reps, torch, flush=None):
    """Mean device milliseconds of fn() over `reps` calls (CUDA events on the current stream)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()                    # > L2: the next call starts cold
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return statistics.mean(ms)


def extra_legs(wl, torch, dev, fmap, labels, plan, peak):
    import ctypes as Cc
    import cases
    legs = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)            # 256 MB > 126 MB L2
    B = fmap.shape[0]
    npx = B * H * W
    L = wl.load()
    check = wl._lib.check

    def st():
        return Cc.c_void_p(torch.cuda.current_stream().cuda_stream)
    # -- the loss alone on the resident 513^2 batch (169 B/pixel) and the E-step + loss sequence
    grad = torch.empty_like(fmap)
    work = torch.empty(2, dtype=torch.float64, device=dev)
    loss = torch.empty((), dtype=torch.float32, device=dev)
    valid = torch.empty((), dtype=torch.int32, device=dev)

    def ce_only():
        check(L.weaklabel_softmax_ce_device(fmap.data_ptr(), labels.data_ptr(), npx, C, 255, loss.data_ptr(),
                                            valid.data_ptr(), grad.data_ptr(), work.data_ptr(), st()))

    def estep_ce():
        check(L.weaklabel_estep_ce_device(plan.handle, fmap.data_ptr(), labels.data_ptr(), loss.data_ptr(),
                                          valid.data_ptr(), grad.data_ptr(), work.data_ptr(), st()))
    ms = timed(ce_only, 5, torch)
    legs["softmax_ce_513x256"] = dict(ms=ms, gbs=npx * CE_BYTES_PER_PIXEL / ms / 1e6,
                                      frac=npx * CE_BYTES_PER_PIXEL / ms / 1e6 / peak,
                                      alg_bytes_per_pixel=CE_BYTES_PER_PIXEL,
                                      kernels="ce_count_kernel + ce_fused_kernel (external labels)",
                                      note="inputs 5.7 GB >> L2")
    ms = timed(estep_ce, 5, torch)
    both = ALG_BYTES_PER_PIXEL + CE_BYTES_PER_PIXEL
    legs["estep_plus_ce_513x256"] = dict(ms=ms, gpixel_s=npx / ms / 1e6, gbs=npx * both / ms / 1e6,
                                         frac=npx * both / ms / 1e6 / peak, alg_bytes_per_pixel=both,
                                         kernels="estep_onchip_kernel + ce_fused_kernel (2 launches, denominator known)")
    del grad
    # -- the in-place map (weaklabel.cpp:108-112) on a copy of the batch (168 B/pixel)
    fcopy = fmap.clone()
    plan.set_outputs(sums=True)
    ms_sums = timed(lambda: plan.labels(fmap, out=labels), 3, torch)
    legs["estep_with_sums_513x256"] = dict(ms=ms_sums, gpixel_s=npx / ms_sums / 1e6, frac=npx * ALG_BYTES_PER_PIXEL / ms_sums / 1e6 / peak,
                                           note="labels + mean_max_0 / mean_max_1 of every image (what the in-place map needs)")
    ms = timed(lambda: plan.finalize_map(fcopy), 3, torch)
    plan.set_outputs(sums=False)
    legs["finalize_map_513x256"] = dict(ms=ms, gbs=npx * MAP_BYTES_PER_PIXEL / ms / 1e6,
                                        frac=npx * MAP_BYTES_PER_PIXEL / ms / 1e6 / peak,
                                        alg_bytes_per_pixel=MAP_BYTES_PER_PIXEL, kernels="finalize_tile_kernel")
    del fcopy
    # -- EM-Fixed labels (fixed biases, no selection chain): a pure 85 B/pixel stream
    sets_b = label_sets(B)
    masks = torch.tensor([sum(1 << k for k in sset) for sset in sets_b], dtype=torch.int32, device=dev)
    ms = timed(lambda: wl.fixed_bias_labels(fmap, masks, 0.5, -0.25), 5, torch)
    legs["em_fixed_labels_513x256"] = dict(ms=ms, gpixel_s=npx / ms / 1e6, gbs=npx * ALG_BYTES_PER_PIXEL / ms / 1e6,
                                           frac=npx * ALG_BYTES_PER_PIXEL / ms / 1e6 / peak, alg_bytes_per_pixel=ALG_BYTES_PER_PIXEL,
                                           kernels="fixed_bias_tile_kernel")
    # -- worst-case label sets at 513^2: quotas that add up to the whole image (P = 4: every step selects), P = 21
    for name, sset, nimg in (("worst_P4_513", [0, 3, 8, 15], 60), ("worst_P21_513", list(range(21)), 30)):
        nimg = min(nimg, B)
        pl = wl.EStepPlan(nimg, H, W, C, device=dev.index)
        pl.prepare(presence_matrix([sset] * nimg), *PARGS, rng=wl.OrderStream(1))
        out = torch.empty((nimg, H, W), dtype=torch.uint8, device=dev)
        ms = timed(lambda: pl.labels(fmap[:nimg], out=out), 3, torch)
        legs[name] = dict(images=nimg, ms=ms, gpixel_s=nimg * H * W / ms / 1e6,
                          frac=nimg * H * W * ALG_BYTES_PER_PIXEL / ms / 1e6 / peak, status=pl.status())
        pl.close()
    # -- config 2: 30 x 41 x 41 x 21 (the shape the reference trains on, deeplab.py:288-293): us per launch
    f2, l2, _ = cases.build_case("c2_41_b30")
    t2 = torch.from_numpy(np.ascontiguousarray(f2)).to(dev)
    pl2 = wl.EStepPlan(*f2.shape, device=dev.index)
    present2 = wl.label_set(l2, C)
    pl2.prepare(present2, *PARGS, rng=wl.OrderStream(1))
    out2 = torch.empty(f2.shape[:3], dtype=torch.uint8, device=dev)
    ms = timed(lambda: pl2.labels(t2, out=out2), 20, torch, flush=flush)
    t0 = time.perf_counter()
    for _ in range(20):
        pl2.prepare(present2, *PARGS, rng=wl.OrderStream(1))
    prep_us = (time.perf_counter() - t0) / 20 * 1e6
    gt2 = torch.from_numpy(np.ascontiguousarray(l2)).to(dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        pl2.prepare_from_gt(gt2, *PARGS, rng=wl.OrderStream(1))
    prep_gt_us = (time.perf_counter() - t0) / 20 * 1e6
    f2c = np.ascontiguousarray(f2)
    wl.estep_c(f2c.copy(), l2, *PARGS)
    t0 = time.perf_counter()
    for _ in range(10):
        wl.estep_c(f2c, l2, *PARGS)
    leg_us = (time.perf_counter() - t0) / 10 * 1e6
    n2 = f2.shape[0] * 41 * 41
    legs["config2_30x41x41"] = dict(labels_us_per_launch=ms * 1e3, gpixel_s=n2 / ms / 1e6,
                                    prepare_host_us=prep_us, prepare_from_device_gt_us=prep_gt_us,
                                    legacy_e_step_us_per_call=leg_us, legacy_gpixel_s=n2 / leg_us / 1e3,
                                    note="latency-bound regime: 30 images on 30 clusters, L2 flushed between launches; "
                                         "legacy = host numpy buffers through e_step (H2D + kernels + D2H)")
    pl2.close()
    # -- config 3: 30 x 321 x 321 x 21 with the fused loss (E-step + CE, 85 + 169 B/pixel)
    f3, l3, _ = cases.build_case("c3_321_b30")
    t3 = torch.from_numpy(np.ascontiguousarray(f3)).to(dev)
    pl3 = wl.EStepPlan(*f3.shape, device=dev.index)
    pl3.prepare(wl.label_set(l3, C), *PARGS, rng=wl.OrderStream(1))
    out3 = torch.empty(f3.shape[:3], dtype=torch.uint8, device=dev)
    g3 = torch.empty_like(t3)
    n3 = f3.shape[0] * 321 * 321

    def c3():
        check(L.weaklabel_estep_ce_device(pl3.handle, t3.data_ptr(), out3.data_ptr(), loss.data_ptr(), valid.data_ptr(),
                                          g3.data_ptr(), work.data_ptr(), st()))
    ms_both = timed(c3, 10, torch, flush=flush)
    ms_lab = timed(lambda: pl3.labels(t3, out=out3), 10, torch, flush=flush)
    legs["config3_30x321x321"] = dict(estep_plus_ce_ms=ms_both, labels_ms=ms_lab, gpixel_s=n3 / ms_both / 1e6,
                                      gbs=n3 * both / ms_both / 1e6, frac=n3 * both / ms_both / 1e6 / peak,
                                      alg_bytes_per_pixel=both, status=pl3.status(),
                                      note="260 MB of scores, L2 flushed between launches")
    # -- the same config with the scores = the 41 x 41 fc8 tensor up-sampled (network.py:40).  Fused: the
    #    producers interpolate the present classes, nothing of 84 B/pixel is ever written or read; against up-sampling
    #    into HBM first (the materialising kernel + the ordinary E-step).  Its own denominator: no HBM roofline is claimed.
    fc3 = torch.randn((f3.shape[0], 41, 41, C), device=dev, generator=torch.Generator(device=dev).manual_seed(5)) * 3.0
    up3 = torch.empty_like(t3)
    ms_fused = timed(lambda: pl3.labels_upsampled(fc3, out=out3), 10, torch, flush=flush)
    fused_lab = out3.clone()

    def mat3():
        wl.upsample_bilinear(fc3, 321, 321, out=up3)
        pl3.labels(up3, out=out3)
    ms_mat = timed(mat3, 10, torch, flush=flush)
    ms_up = timed(lambda: wl.upsample_bilinear(fc3, 321, 321, out=up3), 10, torch, flush=flush)
    legs["config3_fused_upsample_41_to_321"] = dict(
        fused_labels_ms=ms_fused, upsample_then_labels_ms=ms_mat, upsample_alone_ms=ms_up,
        upsample_write_gbs=n3 * 4 * C / ms_up / 1e6, gpixel_s=n3 / ms_fused / 1e6,
        identical_labels=bool(torch.equal(fused_lab, out3)), status=pl3.status(),
        note="labels of the E-step on resize_bilinear(fc8 41x41 -> 321x321): interpolated inside the producers vs "
             "materialised first (84 B/pixel written + read)")
    pl3.close()
    # -- the same at the metric's shape: 256 x 513 x 513 from 65 x 65 (the resident plan and its label sets)
    fc5 = torch.randn((B, 65, 65, C), device=dev, generator=torch.Generator(device=dev).manual_seed(6)) * 3.0
    ms_f5 = timed(lambda: plan.labels_upsampled(fc5, out=labels), 5, torch)
    legs["fused_upsample_65_to_513x256"] = dict(fused_labels_ms=ms_f5, gpixel_s=npx / ms_f5 / 1e6, status=plan.status(),
                                                note="no 84 B/pixel stream at all: bound by the producers' interpolation and the selection chain")
    return legs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--images", type=int, default=IMAGES_PER_GPU, help="images per GPU")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-sample", type=int, default=0, help="images in the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-legs", action="store_true", help="skip the extra configurations / kernels")
    ap.add_argument("--parity-images", type=int, default=16)
    ap.add_argument("--tune", default="", help="cluster,tile,clusters for weaklabel_plan_tune")
    ap.add_argument("--path", default="auto", help="auto | generic | fast")
    ap.add_argument("--phase-profile", default="", help="write per-image phase stamps of one extra run to this .npy")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries ONE JSON line: whatever a library prints there (NCCL's version banner / INFO log) is sent to
    # stderr instead -- fd 1 is re-pointed at fd 2 and the line goes out through a private copy of the real stdout
    global REAL_STDOUT
    if REAL_STDOUT is None:
        sys.stdout.flush()
        REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun exactly as the driver would
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd, stdout=REAL_STDOUT)

    if args.impl == "reference":
        return run_reference_arm(args, rank)

    # ---- CPU baseline first (rank 0, N=1 only), before CUDA is initialised in this process ----
    n_total = args.images * world
    sets = label_sets(n_total)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = len(os.sched_getaffinity(0))
        n_cpu = args.cpu_sample or max(cores, min(64, 4 * cores))
        info = cpu_reference_pass(strided_sample(n_cpu, n_total), sets, cores)
        cpu = dict(value=n_cpu * H * W / info["wall_s"] / 1e9, unit="Gpixel/s", cores=info["cores"],
                   kind=info["kind"],
                   sample="%d 513x513x21 images striding through the run's 256 label sets, one image per e_step "
                          "call (weaklabel.cpp compiled -O2, unmodified), %d worker processes, %.1f s wall" % (
                              n_cpu, info["cores"], info["wall_s"]))
        if not args.no_legs:
            cpu["variants"] = cpu_variants(sets, cores)

    import torch
    import em_adapt_tensorflow_b200 as wl
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (libweaklabel_cuda has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    lo, hi = wl.shard_range(n_total, world, rank)
    B = hi - lo
    present_all = presence_matrix(sets)
    present = present_all[lo:hi]
    fmap = device_scores(lo, hi, dev)
    labels = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
    plan = wl.EStepPlan(B, H, W, C, device=local_rank)
    if args.tune:
        plan.tune(*[int(x) for x in args.tune.split(",")])
    if args.path != "auto":
        plan.set_path(args.path)
    plan.set_outputs(sums=False)             # the metric's path: scores -> labels (deeplab.py:120-123 feeds the loss)
    draws_lo = wl.draws_before(present_all, PARAMS["num_iter"], lo)

    def stream_at_lo():
        s = wl.OrderStream(1)
        s.skip(draws_lo)
        return s
    plan.prepare(present, *PARGS, rng=stream_at_lo())

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        plan.labels(fmap, out=labels)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = plan.launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for s in range(args.steps):
        plan.labels(fmap, out=labels)
        ev[s + 1].record()
    barrier()
    launches = plan.launches - l0
    status = plan.status()                   # NaN / lost-arrival codes of all timed launches (0 = fine)
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch_ms = [ev[s].elapsed_time(ev[s + 1]) for s in range(args.steps)]
    clocks = sampler.result()
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    max_ms = float(t.item())
    # per-rank kernel time and clocks (names the straggler of a multi-GPU run)
    mine = torch.tensor([statistics.mean(per_launch_ms), float(clocks["sm_mhz"] or 0)], dtype=torch.float64, device=dev)
    per_rank = [mine.clone() for _ in range(world)]
    if dist is not None:
        dist.all_gather(per_rank, mine)
    per_rank = [dict(rank=r, kernel_ms=float(v[0].item()), sm_mhz=float(v[1].item())) for r, v in enumerate(per_rank)]

    # ---- the per-step host cost a training loop pays: prepare() (label sets -> class order -> async upload) ----
    t0 = time.perf_counter()
    for _ in range(5):
        plan.prepare(present, *PARGS, rng=stream_at_lo())
    prepare_ms = (time.perf_counter() - t0) / 5 * 1e3
    pev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    pev[0].record()
    for _ in range(3):
        plan.prepare(present, *PARGS, rng=stream_at_lo())
        plan.labels(fmap, out=labels)
    pev[1].record()
    torch.cuda.synchronize()
    step_with_prepare_ms = pev[0].elapsed_time(pev[1]) / 3

    if args.phase_profile and rank == 0:
        plan.profile(True)
        plan.labels(fmap, out=labels)
        torch.cuda.synchronize()
        np.save(args.phase_profile, plan.profile(False, read=True))

    # ---- parity gates: (1) the first images of the shard against the oracle (checker only);
    #      (2) ALL images: the generic kernel (a different algorithm) on the same resident map ----
    parity = None
    if args.parity_images > 0:
        from oracle.oracle import Oracle
        n_chk = min(args.parity_images, B)
        orc = Oracle(1)
        for _ in range(draws_lo):
            orc.rand()
        f_h = fmap[:n_chk].cpu().numpy()
        gt = np.zeros((n_chk, H, W), np.int32)
        for b in range(n_chk):              # any GT map whose stride-2 scan gives the label set
            gt[b].flat[0:2 * len(sets[lo + b]):2] = sets[lo + b]
        want = orc.argmax_labels(orc.e_step(f_h, gt, **PARAMS))
        got = labels[:n_chk].cpu().numpy()
        parity = dict(images=n_chk, mismatched_pixels=int((want != got).sum()))
        if parity["mismatched_pixels"]:
            raise SystemExit("bench.py: labels differ from the oracle: %r" % parity)
        if plan.last_path == "fast":
            alt = wl.EStepPlan(B, H, W, C, device=local_rank)
            alt.set_path("generic")
            alt.prepare(present, *PARGS, rng=stream_at_lo())
            other = alt.labels(fmap)
            same = bool(torch.equal(other, labels))
            parity["all_images_vs_generic_kernel"] = dict(images=B, identical=same, status=alt.status())
            alt.close()
            del other
            if not same:
                raise SystemExit("bench.py: the on-chip and the generic kernel disagree on the full batch")

    # ---- e2e: legacy C-ABI e_step on HOST buffers (H2D + kernels + D2H every step), pinned and pageable ----
    e2e = None
    if not args.no_e2e:
        import cases
        host = torch.empty((B, H, W, C), dtype=torch.float32, pin_memory=True)
        host.copy_(fmap)
        f_np = host.numpy()
        gt = np.stack([cases.paint_label_map(H, W, sets[lo + b], np.random.default_rng(99 + lo + b))
                       for b in range(B)]).astype(np.int32)
        assert np.array_equal(wl.label_set(gt, C), present)
        torch.cuda.synchronize()
        wl.estep_c(f_np, gt, *PARGS)
        host.copy_(fmap)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            wl.estep_c(f_np, gt, *PARGS)
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / args.e2e_steps], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        nbytes = B * H * W * C * 4
        e2e = dict(value=n_total * H * W / float(dt.item()) / 1e9, unit="Gpixel/s",
                   h2d_bytes_per_step=nbytes, d2h_bytes_per_step=nbytes,
                   api="e_step(map, shape, stride, label, ...) legacy C-ABI, in place on pinned host memory; blocks of "
                       "images pipelined on three streams (H2D || E-step + map || D2H)",
                   steps=args.e2e_steps, link_gbs_each_way=nbytes / float(dt.item()) / 1e9)
        del host, f_np
        if world == 1:
            f_pg = fmap.cpu().numpy()            # pageable: what estep.py:21-22 hands to the library
            wl.estep_c(f_pg, gt, *PARGS)
            f_pg = fmap.cpu().numpy()
            t0 = time.perf_counter()
            wl.estep_c(f_pg, gt, *PARGS)
            e2e["pageable"] = dict(value=B * H * W / (time.perf_counter() - t0) / 1e9, unit="Gpixel/s",
                                   note="pageable numpy buffer, staged through the library's pinned ring by host threads")
            del f_pg

    legs = None
    if rank == 0 and world == 1 and not args.no_legs:
        peak0, _ = measured_peak()
        legs = extra_legs(wl, torch, dev, fmap, labels, plan, peak0)

    if rank == 0:
        peak, peak_src = measured_peak()
        k_ms = statistics.mean(per_launch_ms)
        achieved = B * H * W * ALG_BYTES_PER_PIXEL / (k_ms * 1e-3) / 1e9
        kernel = {"fast": "estep_onchip_kernel", "generic": "estep_cluster_kernel"}.get(plan.last_path, "?")
        traffic, traffic_src = ncu_traffic(kernel, B)
        line = dict(metric=METRIC, value=n_total * H * W * args.steps / (max_ms * 1e-3) / 1e9, unit="Gpixel/s",
                    n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=max_ms / args.steps,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                    config=dict(workload="256x513x513x21 E-step per GPU (BASELINE.json configs[3])",
                                images_per_gpu=B, H=H, W=W, C=C, suppress_others=1, num_iter=5,
                                margin_others=1e-5, bg=0.4, fg=0.2, label_sets="VOC-style bg+0..3 fg",
                                outputs="labels (uint8); re-centring sums formed only where a label can depend on them",
                                l2="inputs 5.66 GB per GPU >> 126 MB L2 (no flush needed)",
                                parallelism="image-sharded x%d, no collective" % world),
                    roofline=dict(bound="hbm", achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak,
                                  traffic=traffic, traffic_source=traffic_src, kernel=kernel,
                                  alg_bytes_per_pixel=ALG_BYTES_PER_PIXEL, kernel_ms=k_ms, peak_source=peak_src),
                    cpu_baseline=cpu, e2e=e2e, gpu_launches=int(launches), clocks=clocks, parity=parity,
                    kernel_status=status,
                    prepare=dict(host_ms=prepare_ms, step_with_prepare_ms=step_with_prepare_ms,
                                 note="prepare() = label-set validation + class order from rand() on the host + one "
                                      "asynchronous table upload; it overlaps the previous step's kernel"),
                    per_rank=per_rank, legs=legs)
        print(json.dumps(line), file=REAL_STDOUT or sys.stdout, flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
