"""Experiment: what PCIe gives for the host step's copies — one block vs ranges of rows
([8][B] obs as 2-D blocks), one direction vs both at once."""
import os
import sys
import time

import torch
from cuda.bindings import runtime as rt

B = 1 << 20
K = rt.cudaMemcpyKind
d_obs = torch.zeros((8, B), device="cuda"); d_rew = torch.zeros(B, device="cuda"); d_done = torch.zeros(B, dtype=torch.uint8, device="cuda")
d_act = torch.zeros((2, B), device="cuda")
h_obs = torch.zeros((8, B)).pin_memory(); h_rew = torch.zeros(B).pin_memory(); h_done = torch.zeros(B, dtype=torch.uint8).pin_memory()
h_act = torch.zeros((2, B)).pin_memory()
s_in = torch.cuda.Stream(); s_out = torch.cuda.Stream()


ROWS = "--rows" in sys.argv


def d2h(C, stream):
    L = B // C
    for c in range(C):
        off = c * L
        if ROWS:
            for r in range(8):
                rt.cudaMemcpyAsync(h_obs.data_ptr() + 4 * (r * B + off), d_obs.data_ptr() + 4 * (r * B + off), 4 * L,
                                   K.cudaMemcpyDeviceToHost, stream.cuda_stream)
        else:
            rt.cudaMemcpy2DAsync(h_obs.data_ptr() + 4 * off, 4 * B, d_obs.data_ptr() + 4 * off, 4 * B, 4 * L, 8,
                                 K.cudaMemcpyDeviceToHost, stream.cuda_stream)
        rt.cudaMemcpyAsync(h_rew.data_ptr() + 4 * off, d_rew.data_ptr() + 4 * off, 4 * L, K.cudaMemcpyDeviceToHost,
                           stream.cuda_stream)
        rt.cudaMemcpyAsync(h_done.data_ptr() + off, d_done.data_ptr() + off, L, K.cudaMemcpyDeviceToHost,
                           stream.cuda_stream)


def h2d(C, stream):
    L = B // C
    for c in range(C):
        off = c * L
        if ROWS:
            for r in range(2):
                rt.cudaMemcpyAsync(d_act.data_ptr() + 4 * (r * B + off), h_act.data_ptr() + 4 * (r * B + off), 4 * L,
                                   K.cudaMemcpyHostToDevice, stream.cuda_stream)
        else:
            rt.cudaMemcpy2DAsync(d_act.data_ptr() + 4 * off, 4 * B, h_act.data_ptr() + 4 * off, 4 * B, 4 * L, 2,
                                 K.cudaMemcpyHostToDevice, stream.cuda_stream)


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


for C in (1, 2, 4, 8, 16):
    a = timeit(lambda: d2h(C, s_out))
    b = timeit(lambda: h2d(C, s_in))
    c = timeit(lambda: (h2d(C, s_in), d2h(C, s_out)))
    print(f"C={C:2d}: D2H {a:.3f} ms ({37 * B / a / 1e6:.1f} GB/s)  H2D {b:.3f} ms  both at once {c:.3f} ms", flush=True)
