"""cotix_narrowphase_static — the wavefront sweep with its scratch in a file-scope __device__ array
(no cudaMalloc in the call, no counterpart in the
reference).  Collected LAST (file name) so that a failure here cannot hide another test under -x.
Checked through the C-ABI against the existing caller-workspace entry point cotix_narrowphase_ws:
bit-identical (the tolerance of every narrowphase test here), on two calls in a row that share the
array, plus the size limit.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from gen_inputs import random_pairs

pytestmark = pytest.mark.gpu


def _eq(a, b, what):
    a, b = a.cpu().numpy(), b.cpu().numpy()
    ok = (a == b) | (np.isnan(a) & np.isnan(b)) if a.dtype.kind == "f" else (a == b)
    assert ok.all(), f"{what}: {np.count_nonzero(~ok)} mismatches"


def test_static_scratch_sweep_matches_workspace_sweep(cuda_lib):
    import torch

    import parallax_b200

    L = cuda_lib
    prm = parallax_b200._lib.default_narrow_params()
    MAX = 131072  # COTIX_NARROWPHASE_STATIC_MAX_PAIRS
    rng = np.random.default_rng(2024)
    stream = torch.cuda.current_stream().cuda_stream

    def outputs(n):
        return (torch.full((n,), 7, dtype=torch.uint8, device="cuda"), torch.full((n, 2), -3.0, device="cuda"),
                torch.full((n, 2), -3.0, device="cuda"), torch.full((n,), 7, dtype=torch.uint8, device="cuda"))

    # two calls in a row on the shared array: a full-size one, then a smaller ragged one whose queues
    # are shorter than what the first call left behind
    for n in (MAX, 50001):
        a, b = random_pairs(rng, n)
        da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        nbytes = int(L.cotix_narrowphase_workspace_bytes(n))
        ws = torch.empty(nbytes + 256, dtype=torch.uint8, device="cuda")
        wp = (ws.data_ptr() + 255) & ~255
        ref = outputs(n)
        rc = L.cotix_narrowphase_ws(stream, n, da.data_ptr(), db.data_ptr(), ref[0].data_ptr(), ref[1].data_ptr(),
                                    ref[2].data_ptr(), ref[3].data_ptr(), C.byref(prm), wp, ws.numel() - (wp - ws.data_ptr()))
        assert rc == 0, L.cotix_last_error()
        got = outputs(n)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = L.cotix_narrowphase_static(stream, n, da.data_ptr(), db.data_ptr(), got[0].data_ptr(), got[1].data_ptr(),
                                        got[2].data_ptr(), got[3].data_ptr(), C.byref(prm))
        assert rc == 0, L.cotix_last_error()
        torch.cuda.synchronize()
        if n != MAX:  # (the first call on a device may load the module lazily)
            assert torch.cuda.mem_get_info()[0] >= free0, "the static entry point took device memory"
        for g, r, what in zip(got, ref, ("exists", "penetration", "contact point", "status")):
            _eq(g, r, f"{what} (n = {n})")
        assert 0.2 < float(ref[0].float().mean()) < 0.9  # the batch does contain hits and misses

    # beyond the maximum: a clear error, nothing enqueued
    n = MAX + 1
    a, b = random_pairs(rng, 16)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    rc = L.cotix_narrowphase_static(stream, n, da.data_ptr(), db.data_ptr(), None, None, None, None, C.byref(prm))
    assert rc == -1
    msg = L.cotix_last_error().decode()
    assert "COTIX_NARROWPHASE_STATIC_MAX_PAIRS" in msg and str(MAX) in msg
    assert L.cotix_narrowphase_static(stream, 0, None, None, None, None, None, None, C.byref(prm)) == 0
