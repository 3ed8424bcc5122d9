#!/usr/bin/env python
"""Raw host<->device copy bandwidth of the box with pinned memory: each direction alone and both at once
(what bounds the legacy e_step ABI, which moves 84 B/pixel in and 84 B/pixel out)."""
import torch
n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, reps=5):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    b.record(); torch.cuda.synchronize()
    return reps * n / a.elapsed_time(b) / 1e6
run(True, True, 2)
print("H2D alone %.1f GB/s  D2H alone %.1f GB/s  both at once %.1f GB/s each way" % (run(True, False), run(False, True), run(True, True)))
