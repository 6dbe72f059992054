"""Experiment: PCIe copy rate vs size and direction (1-D copies, back to back on one stream)."""
import time
import torch
from cuda.bindings import runtime as rt

K = rt.cudaMemcpyKind
N = 64 << 20
d = torch.zeros(N, dtype=torch.uint8, device="cuda")
h = torch.zeros(N, dtype=torch.uint8).pin_memory()
s = torch.cuda.Stream()
for size in (64 << 10, 256 << 10, 512 << 10, 1 << 20, 2 << 20, 4 << 20, 8 << 20, 32 << 20):
    out = []
    for kind, dst, src in ((K.cudaMemcpyHostToDevice, d, h), (K.cudaMemcpyDeviceToHost, h, d)):
        reps = max(4, min(64, N // size))
        for trial in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for r in range(reps):
                off = (r * size) % (N - size + 1)
                rt.cudaMemcpyAsync(dst.data_ptr() + off, src.data_ptr() + off, size, kind, s.cuda_stream)
            t1 = time.perf_counter()
            torch.cuda.synchronize()
            t2 = time.perf_counter()
        out.append(f"{size * reps / (t2 - t0) / 1e9:6.1f} GB/s ({(t2 - t0) / reps * 1e6:7.1f} us/copy, issue {(t1 - t0) / reps * 1e6:5.1f} us)")
    print(f"{size >> 10:6d} KiB  H2D {out[0]}   D2H {out[1]}", flush=True)
