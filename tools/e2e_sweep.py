import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
import bench, cases, em_adapt_tensorflow_b200 as wl
B = 256
sets = bench.label_sets(B)
dev = torch.device("cuda", 0)
fmap = bench.device_scores(0, B, dev)
host = torch.empty((B, 513, 513, 21), dtype=torch.float32, pin_memory=True); host.copy_(fmap)
gt = np.stack([cases.paint_label_map(513, 513, sets[b], np.random.default_rng(99 + b)) for b in range(B)]).astype(np.int32)
f = host.numpy()
t0 = time.perf_counter(); p = wl.label_set(gt, 21); print("host label_set single thread: %.1f ms" % ((time.perf_counter() - t0) * 1e3))
for mb in (96, 32, 192, 48):
    os.environ["WEAKLABEL_BLOCK_MB"] = str(mb)
    wl.estep_c(f, gt, *bench.PARGS)
    t0 = time.perf_counter()
    for _ in range(2): wl.estep_c(f, gt, *bench.PARGS)
    dt = (time.perf_counter() - t0) / 2
    print("block %3d MB: %.1f ms per call, %.3f Gpixel/s, %.1f GB/s each way" % (mb, dt * 1e3, B * 513 * 513 / dt / 1e9, f.nbytes / dt / 1e9), flush=True)
