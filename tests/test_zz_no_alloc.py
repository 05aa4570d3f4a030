"""gb200_second_peak_noalloc (no counterpart in the reference): the second-peak search
with its scratch in a fixed device array, for callers that cannot call cudaMalloc.  Checked through the C ABI against the existing
entry point gb200_acq_second_peak on the same kept-grid rows, on calls in a row that share the scratch (a long row, then a short
one, then the long one again), plus the size limit.  This file is collected last on purpose."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [HERE]
import synth  # noqa: E402

pytestmark = pytest.mark.gpu


def test_second_peak_noalloc_matches_existing_entry_point(service):
    import torch

    import gnss_b200 as gb

    lib = gb.lib()
    fs = 4000000
    sats = [dict(kind="gps_l1_ca", prn=5, cn0_dbhz=50.0, doppler_hz=1250.0, delay_samples=1234)]
    x = synth.make_iq(fs, 8000, sats, seed=77)
    service.push(3, x, gb.ITEM_32FC, 0)
    cases = []
    # a 2 ms (8000-point) and a 1 ms (4000-point) search, each keeping its magnitude grid
    for ms in (2, 1):
        n = 4000 * ms
        code = synth.sampled_code(synth.code_table("gps_l1_ca", 5), 1.023e6, fs, n)
        a = gb.Acquisition(service, sampled_ms=ms, ms_per_code=ms, doppler_max=2500, doppler_step=250, threshold=1e-9)
        a.keep_grid(True)
        a.set_local_code(code)
        r = a.run(3, 0)
        grid = a.read_grid()
        assert grid.shape[1] == n
        for bin_, idx, hw in ((r.doppler_bin, r.code_phase_index, 4), (0, 0, 7), (grid.shape[0] - 1, n - 1, 4), (3, n // 2, 1)):
            want = C.c_float()
            service.check(lib.gb200_acq_second_peak(a.h, bin_, idx, hw, C.byref(want)))
            cases.append((grid[bin_].copy(), idx, hw, want.value))
        a.close()
    stream = torch.cuda.current_stream().cuda_stream
    err = C.create_string_buffer(256)
    out = torch.zeros(len(cases) * 2, dtype=torch.float32, device="cuda")
    rows = [torch.from_numpy(c[0]).cuda() for c in cases]
    torch.cuda.synchronize()
    # calls in a row on one stream, alternating long and short rows so every call meets the previous call's contents in the scratch
    order = [0, 4, 1, 5, 2, 6, 3, 7] + list(range(8))
    for slot, k in enumerate(order):
        row, idx, hw, _ = cases[k]
        rc = lib.gb200_second_peak_noalloc(rows[k].data_ptr(), len(row), idx, hw, out.data_ptr() + 4 * slot, stream, err, len(err))
        assert rc == 0, err.value
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for slot, k in enumerate(order):
        want = cases[k][3]
        assert abs(got[slot] - want) <= 1e-4 * abs(want), (slot, k, got[slot], want)
        assert got[slot] == want  # the same kernel arithmetic: equal, not just close
    # the detected cell's row really had its main peak excluded
    assert cases[0][3] < cases[0][0].max()
    # limit and argument checks: refused with a message, nothing launched
    big = torch.zeros(gb_max_row(lib) + 1, dtype=torch.float32, device="cuda")
    rc = lib.gb200_second_peak_noalloc(big.data_ptr(), big.numel(), 0, 4, out.data_ptr(), stream, err, len(err))
    assert rc == -4 and b"GB200_NOALLOC_MAX_ROW" in err.value
    assert lib.gb200_second_peak_noalloc(rows[0].data_ptr(), 8000, 8000, 4, out.data_ptr(), stream, err, len(err)) == -1
    assert lib.gb200_second_peak_noalloc(None, 8000, 0, 4, out.data_ptr(), stream, None, 0) == -1
    # a row of exactly the maximum length is served
    rng = np.random.default_rng(3)
    full = rng.random(gb_max_row(lib)).astype(np.float32)
    dfull = torch.from_numpy(full).cuda()
    assert lib.gb200_second_peak_noalloc(dfull.data_ptr(), len(full), 100, 50, out.data_ptr(), stream, err, len(err)) == 0, err.value
    torch.cuda.synchronize()
    ref = full.copy()
    ref[50:150] = 0
    assert out[0].item() == ref.max()


def gb_max_row(lib):
    return 262144  # GB200_NOALLOC_MAX_ROW (include/gnss_b200.h)
