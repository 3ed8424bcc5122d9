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
legacy C-ABI `e_step` on HOST (pinned) buffers, H2D + D2H inside the timed region.
`cpu_baseline` = the reference's own weaklabel.cpp (oracle/_ref, compiled unmodified) on the
host cores over a bounded sample.  `--impl reference` times that CPU implementation as its own
arm.  Only the cpu_baseline / reference legs and the parity gate touch oracle/.
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
ALG_BYTES_PER_PIXEL = 85          # 21*4 read + 1 label byte
METRIC = "E-step Gpixel/s (21-class, 513^2, B200) and % of HBM roofline"


# ------------------------------------------------------------------------------------------
# synthetic workload
# ------------------------------------------------------------------------------------------
def label_sets(n_images, seed=4321):
    import cases
    return cases.voc_label_sets(n_images, np.random.default_rng(seed))


def presence_matrix(sets):
    p = np.zeros((len(sets), C), np.uint8)
    for b, s in enumerate(sets):
        p[b, s] = 1
    return p


def device_scores(lo, hi, device):
    import torch
    out = torch.empty((hi - lo, H, W, C), dtype=torch.float32, device=device)
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
    which, idxs, sets, barrier_t = args
    import cases
    from oracle.oracle import Oracle, Reference
    if which == "reference":
        ref = Reference("ref")
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
        if which == "reference":
            ref.e_step_image_inplace(f, l, **PARAMS)
        else:
            orc.e_step_inplace(f, l, **PARAMS)
    return t0, time.time(), len(data)


def cpu_reference_pass(n_images, sets, cores):
    """All-core wall time of the reference CPU E-step over `n_images` 513^2 images."""
    import multiprocessing as mp
    from oracle.oracle import Reference, build
    build()
    which = "reference" if Reference.available("ref") else "port"
    per = [list(range(k, n_images, cores)) for k in range(cores)]
    per = [p for p in per if p]
    gen_s = 0.9 * max(len(p) for p in per) + 3.0          # generation ~0.6 s/image/core
    start_at = time.time() + gen_s
    ctx = mp.get_context("spawn")
    with ctx.Pool(len(per)) as pool:
        res = pool.map(_cpu_worker, [(which, p, sets, start_at) for p in per])
    t0 = min(r[0] for r in res)
    t1 = max(r[1] for r in res)
    late = max(r[0] for r in res) - start_at
    return dict(kind=which, wall_s=t1 - t0, images=sum(r[2] for r in res), cores=len(per), late_s=late)


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


def ncu_traffic():
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["estep_bytes_per_launch"]
    except Exception:
        return None


def run_reference_arm(args, rank):
    if rank != 0:
        return 0
    cores = len(os.sched_getaffinity(0))
    n_img = max(cores, 2 * cores if cores <= 16 else cores)
    sets = label_sets(max(IMAGES_PER_GPU, n_img))
    times = []
    info = None
    for step in range(args.warmup + args.steps):
        info = cpu_reference_pass(n_img, sets, cores)
        if step >= args.warmup:
            times.append(info["wall_s"])
    pixels = n_img * H * W
    t = sum(times) / len(times)
    val = pixels / t / 1e9
    sample = "%d of the 256 513x513x21 images per step (one image per e_step call, %d worker processes)" % (
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
    print(json.dumps(line), flush=True)
    return 0


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
    ap.add_argument("--parity-images", type=int, default=8)
    ap.add_argument("--tune", default="", help="cluster,tile,clusters for weaklabel_plan_tune")
    ap.add_argument("--path", default="auto", help="auto | generic | fast")
    ap.add_argument("--phase-profile", default="", help="write per-image phase stamps of one extra run to this .npy")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun exactly as the driver would
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)

    if args.impl == "reference":
        return run_reference_arm(args, rank)

    # ---- CPU baseline first (rank 0, N=1 only), before CUDA is initialised in this process ----
    n_total = args.images * world
    sets = label_sets(n_total)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = len(os.sched_getaffinity(0))
        n_cpu = args.cpu_sample or max(cores, min(64, 4 * cores))
        info = cpu_reference_pass(n_cpu, sets, cores)
        cpu = dict(value=n_cpu * H * W / info["wall_s"] / 1e9, unit="Gpixel/s", cores=info["cores"],
                   kind=info["kind"],
                   sample="%d 513x513x21 images of the same synthetic distribution, one image per e_step "
                          "call, %d worker processes, %.1f s wall" % (n_cpu, info["cores"], info["wall_s"]))

    import torch
    import em_adapt_tensorflow_b200 as wl
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (libweaklabel_cuda has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        os.environ.pop("NCCL_DEBUG", None)     # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        os.environ["NCCL_DEBUG_FILE"] = os.devnull
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
    stream = wl.positioned_stream(present_all, PARAMS["num_iter"], lo)
    plan.prepare(present, PARAMS["suppress_others"], PARAMS["num_iter"], PARAMS["margin_others"],
                 PARAMS["bg"], PARAMS["fg"], rng=stream)

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
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch_ms = [ev[s].elapsed_time(ev[s + 1]) for s in range(args.steps)]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    max_ms = float(t.item())

    if args.phase_profile and rank == 0:
        plan.profile(True)
        plan.labels(fmap, out=labels)
        torch.cuda.synchronize()
        np.save(args.phase_profile, plan.profile(False, read=True))

    # ---- parity gate: first images of the shard against the oracle (checker only) ----
    parity = None
    if args.parity_images > 0:
        from oracle.oracle import Oracle
        n_chk = min(args.parity_images, B)
        orc = Oracle(1)
        for _ in range(wl.draws_before(present_all, PARAMS["num_iter"], lo)):
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

    # ---- e2e: legacy C-ABI e_step on pinned HOST buffers (H2D + kernels + D2H every step) ----
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
        wl.estep_c(f_np, gt, *[PARAMS[k] for k in ("suppress_others", "num_iter", "margin_others", "bg", "fg")])
        host.copy_(fmap)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            wl.estep_c(f_np, gt, *[PARAMS[k] for k in ("suppress_others", "num_iter", "margin_others", "bg", "fg")])
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / args.e2e_steps], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        nbytes = B * H * W * C * 4
        e2e = dict(value=n_total * H * W / float(dt.item()) / 1e9, unit="Gpixel/s",
                   h2d_bytes_per_step=nbytes, d2h_bytes_per_step=nbytes,
                   api="e_step(map, shape, stride, label, ...) legacy C-ABI, in place on pinned host memory",
                   steps=args.e2e_steps)
    clocks = sampler.result()

    if rank == 0:
        peak, peak_src = measured_peak()
        k_ms = statistics.mean(per_launch_ms)
        achieved = B * H * W * ALG_BYTES_PER_PIXEL / (k_ms * 1e-3) / 1e9
        line = dict(metric=METRIC, value=n_total * H * W * args.steps / (max_ms * 1e-3) / 1e9, unit="Gpixel/s",
                    n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=max_ms / args.steps,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                    config=dict(workload="256x513x513x21 E-step per GPU (BASELINE.json configs[3])",
                                images_per_gpu=B, H=H, W=W, C=C, suppress_others=1, num_iter=5,
                                margin_others=1e-5, bg=0.4, fg=0.2, label_sets="VOC-style bg+0..3 fg",
                                l2="inputs 5.66 GB per GPU >> 126 MB L2 (no flush needed)",
                                parallelism="image-sharded x%d, no collective" % world),
                    roofline=dict(bound="hbm", achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak,
                                  traffic=ncu_traffic(), kernel={"fast": "estep_onchip_kernel", "generic": "estep_cluster_kernel"}.get(plan.last_path, "?"),
                                  alg_bytes_per_pixel=ALG_BYTES_PER_PIXEL, kernel_ms=k_ms, peak_source=peak_src),
                    cpu_baseline=cpu, e2e=e2e, gpu_launches=int(launches), clocks=clocks, parity=parity)
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
