"""GPU parity of the pcps_acquisition_fine_doppler_cc mirror (B200_Pcps_Acquisition_Fine_Doppler) against the
oracle's restatement of the block: accumulated grid cell and Doppler point, first / second peak and their ratio, the index of the
320000-point zero-padded FFT's maximum and the refined Doppler."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [HERE]
import host_lib as hl  # noqa: E402
import oracle_lib as ol  # noqa: E402
import synth  # noqa: E402

pytestmark = pytest.mark.gpu


def close(a, b, rtol=1e-4):
    return abs(a - b) <= rtol * max(abs(a), abs(b), 1e-30)


@pytest.mark.parametrize("dop,delay,cn0,dwells,thr", [(1730.0, 1234, 50.0, 2, 2.5), (-3120.0, 3999, 47.0, 3, 2.5), (400.0, 0, 50.0, 1, 2.5), (0.0, 100, -100.0, 2, 2.5)])
def test_fine_doppler_block_matches_oracle(service, dop, delay, cn0, dwells, thr):
    import gnss_b200 as gb

    fs, n = 4000000, 4000
    sats = [dict(kind="gps_l1_ca", prn=5, cn0_dbhz=cn0, doppler_hz=dop, delay_samples=delay)] if cn0 > 0 else []
    x = synth.make_iq(fs, 10 * n, sats, seed=int(abs(dop)) + 7)
    code = np.zeros(2 * n, np.float32)
    ol.oracle().go_gps_l1_ca_code_sampled(code, 5, fs, 0)
    code = code.view(np.complex64)
    o = ol.OracleFineDoppler(n, dwells, 5000, 500, fs, thr)
    ro = o.attempt(x, code)
    o.close()
    service.push(3, x, gb.ITEM_32FC, 20000)
    ev, r = hl.fine_doppler_attempt(service, 3, code, 20000, samples_per_ms=n, max_dwells=dwells, doppler_max=5000, doppler_step=500, fs_in=fs, threshold=thr)
    assert ev == (1 if ro.positive else 2), (ev, ro.positive, r)
    if cn0 > 0:
        assert int(r["index_time"]) == ro.index_time and int(r["index_doppler"]) == ro.index_doppler, (r, ro.index_time, ro.index_doppler)
        assert close(r["first_peak"], ro.first_peak) and close(r["second_peak"], ro.second_peak, 1e-3) and close(r["test_statistic"], ro.test_statistic, 1e-3)
        assert r["acq_delay_samples"] == ro.acq_delay_samples
        assert ro.positive == 1 and ro.fine_accepted == 1 and int(r["fine_accepted"]) == 1
        assert int(r["fine_index"]) == ro.fine_index and r["acq_doppler_hz"] == ro.acq_doppler_hz, (r, ro.fine_index, ro.acq_doppler_hz)
        assert abs(r["acq_doppler_hz"] - dop) <= 12.5
    else:
        # noise only: negative like the oracle, no refinement; the peak values agree although the cell is arbitrary
        assert ro.positive == 0 and int(r["fine_accepted"]) == 0
        assert close(r["first_peak"], ro.first_peak, 1e-3)


def test_wiped_spectrum_peak_and_second_peak_entry_points(service):
    """gb200_acq_wiped_spectrum_peak against numpy (a tone times a replica, 8x zero-padded: the peak bin is the tone's), ragged
    lengths, and the error paths of gb200_acq_second_peak."""
    import ctypes as C

    import gnss_b200 as gb

    rng = np.random.default_rng(5)
    n, L = 4000, 32000
    rep = rng.choice([-1.0, 1.0], n).astype(np.complex64)
    k0 = 1234
    x = (rep * np.exp(2j * np.pi * k0 * np.arange(n) / L)).astype(np.complex64)
    service.push(3, x, gb.ITEM_32FC, 0)
    idx, mag = C.c_uint32(), C.c_float()
    lib = gb.lib()
    service.check(lib.gb200_acq_wiped_spectrum_peak(service.h, 3, 0, rep.ctypes.data_as(C.c_void_p), n, L, C.byref(idx), C.byref(mag)))
    spec = np.abs(np.fft.fft(np.concatenate([x * rep, np.zeros(L - n)]))) ** 2
    assert idx.value == int(np.argmax(spec)) == k0 and abs(mag.value - spec.max()) <= 1e-4 * spec.max()
    assert lib.gb200_acq_wiped_spectrum_peak(service.h, 3, 0, rep.ctypes.data_as(C.c_void_p), n, n - 1, C.byref(idx), C.byref(mag)) == -1
    assert lib.gb200_acq_wiped_spectrum_peak(service.h, 3, 10 ** 9, rep.ctypes.data_as(C.c_void_p), n, L, C.byref(idx), C.byref(mag)) == -4
    a = gb.Acquisition(service)
    sec = C.c_float()
    assert lib.gb200_acq_second_peak(a.h, 0, 0, 4, C.byref(sec)) == -5  # keeps no grid
    a.close()
