"""Host<->device copy bandwidth of this box (pinned memory), alone and concurrently.
Explains the end-to-end (host-pointer) numbers of bench.py."""
import time
import torch

n = 256 << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down, reps=8):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    return n * reps / (time.perf_counter() - t0) / 1e9


run(True, True, 2)
print("H2D alone  %.1f GB/s" % run(True, False))
print("D2H alone  %.1f GB/s" % run(False, True))
print("both       %.1f GB/s each direction" % run(True, True))
for mb in (16, 32, 64):
    k = mb << 20
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(0, n, k):
        d_in[i:i + k].copy_(h_in[i:i + k], non_blocking=True)
    torch.cuda.synchronize()
    print("H2D in %d MB pieces %.1f GB/s" % (mb, n / (time.perf_counter() - t0) / 1e9))
