"""GPU parity: PCPS acquisition (libgnss_b200.so via the C ABI) against the restated pcps_acquisition oracle.

Bar: code_phase_index, doppler bin / Hz and decision EQUAL to the oracle on every input where
the float32-DFT oracle and the float64-DFT oracle agree (all inputs here carry a signal or are checked for
that); test_statistic, input_power, peak_mag within 1e-4 relative.
"""
import os

import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-4


def close(a, b, rtol=RTOL):
    return abs(a - b) <= rtol * abs(b)


def check_against_oracle(r, ro, what=""):
    assert (r.code_phase_index, r.doppler_bin, r.doppler_hz, r.decision) == (ro.code_phase_index, ro.doppler_bin, ro.doppler_hz, ro.decision), \
        (what, (r.code_phase_index, r.doppler_bin, r.doppler_hz, r.decision), (ro.code_phase_index, ro.doppler_bin, ro.doppler_hz, ro.decision))
    assert close(r.test_statistic, ro.test_statistic), (what, r.test_statistic, ro.test_statistic)
    assert close(r.peak_mag, ro.peak_mag), (what, r.peak_mag, ro.peak_mag)
    if ro.input_power:
        assert close(r.input_power, ro.input_power), (what, r.input_power, ro.input_power)
    assert r.acq_delay_samples == ro.acq_delay_samples and r.acq_doppler_hz == ro.acq_doppler_hz


def run_pair(service, cfg, code, samples_c64, item="32fc", first=0, stream=2, dwells=1, keep_grid=False):
    """Runs `dwells` consecutive dwells through GPU and oracle; returns list of (gpu, oracle) results."""
    import gnss_b200 as gb

    it = {"32fc": gb.ITEM_32FC, "16ic": gb.ITEM_16IC}[item]
    if item == "32fc":
        service.push(stream, samples_c64, it, first)
        host = samples_c64
    else:
        q = synth.quantize(samples_c64, item)
        service.push(stream, q, it, first)
        host = q
    a = gb.Acquisition(service, item_type=it, **cfg)
    if keep_grid:
        a.keep_grid(True)
    a.set_local_code(code)
    o = ol.OracleAcq(cshort=int(item == "16ic"), keep_grid=int(keep_grid), **cfg)
    o.set_local_code(code)
    assert (a.fft_size, a.consumed, a.num_bins) == (o.fft_size, o.consumed, o.num_bins)
    out = []
    for d in range(dwells):
        off = d * a.consumed
        r = a.run(stream, first + off)
        seg = host[off: off + a.consumed] if item == "32fc" else host[2 * off: 2 * (off + a.consumed)]
        ro = o.run(seg, first + off + a.consumed)
        assert r.acq_samplestamp == ro.acq_samplestamp
        out.append((r, ro))
    grids = (a.read_grid(), o.grid()) if keep_grid else None
    a.close()
    o.close()
    return out, grids


def test_gps_fixture(service, golden_dir):
    """gps_l1_ca_pcps_acquisition_test.cc:285-286, 349-356: PRN 1, delay 524 +-0.5 chip, Doppler 1680 +-666 Hz, positive."""
    x = np.fromfile(golden_dir + "/GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat", np.complex64)
    code = synth.golden_array("gps_prn1_4msps_sampled")
    cfg = dict(doppler_max=5000, doppler_step=100, threshold=0.001, samples_per_chip=4)
    (pairs, grids) = run_pair(service, cfg, code, x, keep_grid=True)
    r, ro = pairs[0]
    check_against_oracle(r, ro, "gps fixture")
    assert r.decision == 1 and abs(r.acq_delay_samples - 524) < 0.5 * 4000 / 1023 and abs(r.acq_doppler_hz - 1680) <= 666
    assert (r.code_phase_index, r.doppler_hz) == (524, 1700)
    g, go = grids
    assert g.shape == go.shape == (100, 4000)
    # whole grid: the engine reproduces the reference's float32-accumulated wipe-off phase exactly (GbWipeSeg), so the
    # grids agree to DFT rounding; against the oracle's exact-phase arbiter they differ by the reference's own phase drift.
    assert np.abs(g - go).max() <= 3e-6 * go.max(), np.abs(g - go).max() / go.max()
    oa = ol.OracleAcq(analytic_wipeoff=1, keep_grid=1, **cfg)
    oa.set_local_code(code)
    oa.run(x[:4000], 4000)
    assert np.abs(g - oa.grid()).max() <= 5e-4 * go.max()


def test_galileo_fixture(service, golden_dir):
    """galileo_e1_pcps_ambiguous_acquisition_test.cc:293-294, 348-358: delay 2920 +-0.175 chip, Doppler -632 +-166 Hz."""
    x = np.fromfile(golden_dir + "/Galileo_E1_ID_1_Fs_4Msps_8ms.dat", np.complex64)
    code = synth.golden_array("gal_e1b_prn1_4msps_sampled")
    cfg = dict(sampled_ms=4, ms_per_code=4, doppler_max=10000, doppler_step=250, threshold=1e-4, samples_per_ms=4000.0,
               samples_per_code=16000.0, samples_per_chip=4)
    (pairs, _) = run_pair(service, cfg, code, x)
    r, ro = pairs[0]
    check_against_oracle(r, ro, "galileo fixture")
    assert r.decision == 1 and r.code_phase_index == 2920 and abs(r.acq_doppler_hz + 632) <= 166


def c1_signal(n, seed=20260001, cn0=58.0):
    """Config C1's constellation at CN0s high enough that every arg-max is well conditioned
    (parity needs separated peaks; detection performance at 45..38 dB-Hz is not what these tests measure)."""
    prns = [1, 3, 6, 9, 12, 17, 22, 28]
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=cn0 - k, doppler_hz=-4310 + 1237 * k, delay_samples=137 + 487 * k) for k, p in enumerate(prns)]
    return prns, synth.make_iq(4e6, n, sats, seed)


def gps_local_code(prn, fs=4000000, n=4000):
    c = np.zeros(2 * n, np.float32)
    ol.oracle().go_gps_l1_ca_code_sampled(c, prn, fs, 0)
    return c.view(np.complex64)


@pytest.mark.parametrize("item", ["32fc", "16ic"])
def test_c1_batch_matches_oracle(service, item):
    """Config C1: 8 present PRNs + 1 absent searched in ONE batch over the same block (shared forward FFTs)."""
    import gnss_b200 as gb

    prns, x = c1_signal(8000)
    it = {"32fc": gb.ITEM_32FC, "16ic": gb.ITEM_16IC}[item]
    host = x if item == "32fc" else synth.quantize(x, item)
    service.push(2, host, it, 5000)
    cfg = dict(doppler_max=5000, doppler_step=250, threshold=0.008, samples_per_chip=4)
    acqs, oracles = [], []
    for p in prns + [31]:
        a = gb.Acquisition(service, item_type=it, **cfg)
        a.set_local_code(gps_local_code(p))
        acqs.append(a)
        o = ol.OracleAcq(cshort=int(item == "16ic"), **cfg)
        o.set_local_code(gps_local_code(p))
        oracles.append(o)
    res = gb.acq_run_batch(acqs, 2, 5000 + 1234)
    for k, (a, o) in enumerate(zip(acqs, oracles)):
        seg = host[1234: 1234 + 4000] if item == "32fc" else host[2 * 1234: 2 * (1234 + 4000)]
        ro = o.run(seg, 5000 + 1234 + 4000)
        if k < 8:
            check_against_oracle(res[k], ro, "prn %d" % prns[k])
            assert res[k].decision == 1
            assert abs(int(res[k].code_phase_index) - (137 + 487 * k - 1234) % 4000) <= 1
        else:  # noise only: decision exact, arg-max arbitrary
            assert res[k].decision == ro.decision == 0
            assert close(res[k].input_power, ro.input_power)
        a.close()
        o.close()


def test_zero_padded_and_bit_transition(service):
    """sampled_ms != ms_per_code -> 2x zero-padded FFT with the code in the second half (:255-269);
    bit_transition_flag -> FFT 4x the coherent length, magnitudes from the second half (:113-117, :730)."""
    prns, x = c1_signal(20000)
    code2 = np.tile(gps_local_code(1), 2)  # adapter replicates the code sampled_ms times
    cfg = dict(sampled_ms=2, ms_per_code=1, doppler_max=5000, doppler_step=250, threshold=0.005, samples_per_chip=4,
               samples_per_ms=4000.0, samples_per_code=4000.0)
    (pairs, _) = run_pair(service, cfg, code2, x)
    check_against_oracle(*pairs[0], "zero padded")
    assert pairs[0][0].decision == 1
    codebt = np.concatenate([gps_local_code(1), np.zeros(4000, np.complex64)])  # [code x T | 0 x T], quirk 10
    cfg = dict(sampled_ms=1, ms_per_code=1, bit_transition_flag=1, doppler_max=5000, doppler_step=250, threshold=0.001,
               samples_per_chip=4, samples_per_ms=4000.0, samples_per_code=4000.0)
    (pairs, _) = run_pair(service, cfg, codebt, x)
    check_against_oracle(*pairs[0], "bit transition")


def test_multi_dwell_and_peak_statistic(service):
    """max_dwells=3 non-coherent accumulation with the first/second-peak statistic (CFAR is disabled for
    max_dwells > 1, :152-159), then use_CFAR_algorithm=false single dwell (E5a default)."""
    prns, x = c1_signal(16000, seed=77)
    cfg = dict(max_dwells=3, doppler_max=5000, doppler_step=500, threshold=1e9, samples_per_chip=4)
    (pairs, _) = run_pair(service, cfg, gps_local_code(28), x, dwells=3)
    for d, (r, ro) in enumerate(pairs):
        check_against_oracle(r, ro, "dwell %d" % d)
    assert [p[0].decision for p in pairs] == [2, 2, 0]
    cfg = dict(use_CFAR_algorithm_flag=0, doppler_max=5000, doppler_step=250, threshold=2.5, samples_per_chip=4)
    (pairs, _) = run_pair(service, cfg, gps_local_code(3), x, dwells=2)
    for d, (r, ro) in enumerate(pairs):
        check_against_oracle(r, ro, "peak stat %d" % d)
    assert pairs[0][0].decision == 1


def test_two_step_acquisition(service):
    """make_two_steps: coarse grid arms step two, fine grid of second_nbins x second_doppler_step confirms (:832-858, :383-390)."""
    prns, x = c1_signal(16000, seed=5)
    cfg = dict(make_2_steps=1, num_doppler_bins_step2=4, doppler_step2=125.0, doppler_max=5000, doppler_step=500, threshold=0.008,
               samples_per_chip=4)
    (pairs, _) = run_pair(service, cfg, gps_local_code(6), x, dwells=2)
    check_against_oracle(*pairs[0], "coarse")
    check_against_oracle(*pairs[1], "fine")
    assert [p[0].decision for p in pairs] == [3, 1]


@pytest.mark.parametrize("n_ms,fs", [(1, 6625000), (1, 2048000), (1, 25000000)])
def test_odd_lengths_random_code(service, n_ms, fs):
    """DFT lengths with a prime factor 53 (NT1065 GLONASS fixture rate), a power of two, and 2^3 5^5 (config C3)."""
    n = fs // 1000 * n_ms
    rng = np.random.default_rng(fs)
    code = rng.choice([-1.0, 1.0], n).astype(np.complex64)
    x = (rng.standard_normal(2 * n) + 1j * rng.standard_normal(2 * n)).astype(np.complex64)
    delay, dop = n // 3, 750.0
    t = np.arange(2 * n) / fs
    x += 0.5 * np.roll(np.tile(code, 2), delay) * np.exp(2j * np.pi * dop * t).astype(np.complex64)
    cfg = dict(doppler_max=2000, doppler_step=250, threshold=0.0005, samples_per_chip=2, fs_in=fs, samples_per_ms=float(fs // 1000),
               samples_per_code=float(n))
    (pairs, _) = run_pair(service, cfg, code, x.astype(np.complex64))
    check_against_oracle(*pairs[0], "n=%d" % n)
    assert pairs[0][0].code_phase_index == delay and pairs[0][0].doppler_hz == 750 and pairs[0][0].decision == 1


@pytest.mark.parametrize("fs", [4000000, 8000000, 16000000, 20000000, 32000000, 40000000, 50000000, 80000000, 200000000])
@pytest.mark.parametrize("keep_grid", [False, True])
def test_fast_inverse_lengths(service, fs, keep_grid):
    """Every (N1, N2) pair with compile-time in-place DIF inverse kernels (csrc/acq_dif.cuh): the fused arg-max path and the
    grid path against the oracle (25000 is covered by test_odd_lengths_random_code)."""
    n = fs // 1000
    rng = np.random.default_rng(fs + 1)
    code = rng.choice([-1.0, 1.0], n).astype(np.complex64)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    delay, dop = (2 * n) // 7, -500.0
    t = np.arange(n) / fs
    x += 0.5 * np.roll(code, delay) * np.exp(2j * np.pi * dop * t).astype(np.complex64)
    cfg = dict(doppler_max=1000, doppler_step=250, threshold=0.0005, samples_per_chip=2, fs_in=fs, samples_per_ms=float(n),
               samples_per_code=float(n))
    (pairs, grids) = run_pair(service, cfg, code, x.astype(np.complex64), keep_grid=keep_grid)
    check_against_oracle(*pairs[0], "n=%d" % n)
    assert pairs[0][0].code_phase_index == delay and pairs[0][0].doppler_hz == -500 and pairs[0][0].decision == 1
    if keep_grid:
        g, go = grids
        assert g.shape == go.shape
        assert np.max(np.abs(g - go)) <= 2e-5 * np.max(go)


def test_bin_range_units_equal_full_grid(service):
    """shard.py's Doppler-bin-range units (sub-grid + freq_offset_hz through the unchanged C ABI) merged with the reference's
    tie-break reproduce the unsharded search exactly (C1 on 8 GPUs shards one PRN's grid by bin)."""
    import gnss_b200 as gb
    import shard

    n = 4000
    x = c1_signal(2 * n, seed=99)[1][:n]
    code = gps_local_code(6)
    service.push(2, x, gb.ITEM_32FC, 0)
    full = gb.Acquisition(service, doppler_max=5000, doppler_step=250, threshold=0.008)
    full.set_local_code(code)
    rf = full.run(2, 0)
    full.close()
    for world in (2, 3, 8):
        units = shard.acquisition_units(1, 5000, 250, world)
        recs = []
        for u, unit in enumerate(units):
            a = gb.Acquisition(service, doppler_max=unit.doppler_max, doppler_step=unit.doppler_step, freq_offset_hz=unit.freq_offset_hz,
                               threshold=0.008)
            assert a.num_bins == unit.bin_hi - unit.bin_lo
            a.set_local_code(code)
            r = a.run(2, 0)
            recs.append(shard.acq_record(u, unit, r.code_phase_index, r.doppler_bin, r.doppler_hz, r.peak_mag, r.input_power))
            a.close()
        m = shard.merge_acq_records(recs, 1, n, threshold=0.008)[0]
        assert (m["code_phase_index"], m["doppler_bin"], m["doppler_hz"], m["decision"]) == (rf.code_phase_index, rf.doppler_bin, rf.doppler_hz, rf.decision)
        assert m["peak_mag"] == rf.peak_mag and abs(m["test_statistic"] - rf.test_statistic) <= 1e-6 * rf.test_statistic


@pytest.mark.parametrize("n", [4000, 32000, 6625])
def test_exact_ties_first_maximum_wins(service, n):
    """All-equal magnitudes (zero input): the reference's scan keeps the FIRST maximum (volk_gnsssdr_32f_index_max_32u, strict '>'
    over bins in ascending order), i.e. code phase 0 of Doppler bin 0 -- on the fast (4000, 32000) and general (6625) inverse paths."""
    import gnss_b200 as gb

    fs = n * 1000
    cfg = dict(doppler_max=1000, doppler_step=250, threshold=0.5, samples_per_chip=2, fs_in=fs, samples_per_ms=float(n), samples_per_code=float(n))
    code = np.zeros(n, np.complex64)
    code[0] = 1.0
    (pairs, _) = run_pair(service, cfg, code, np.zeros(n, np.complex64))
    r, ro = pairs[0]
    assert (r.code_phase_index, r.doppler_bin) == (ro.code_phase_index, ro.doppler_bin) == (0, 0)


def test_tong_detector_block_matches_oracle(service, golden_dir):
    """B200_Pcps_Tong_Acquisition (host mirror of pcps_tong_acquisition_cc on the C ABI: explicit Doppler grid + per-dwell scaled
    accumulation on the device) against the restated block, dwell by dwell: same cell, same counter walk, statistic within 1e-4."""
    import gnss_b200 as gb
    import host_lib as hl

    x = np.fromfile(golden_dir + "/GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat", np.complex64)
    rng = np.random.default_rng(3)
    # four 1 ms dwells: the fixture's two, then the same signal again under fresh noise
    x4 = np.concatenate([x, x + 0.02 * (rng.standard_normal(8000) + 1j * rng.standard_normal(8000)).astype(np.complex64)]).astype(np.complex64)
    service.push(3, x4, gb.ITEM_32FC, 0)
    code = synth.golden_array("gps_prn1_4msps_sampled")[:4000]
    for thr, init, mx, want_event in ((0.3, 1, 4, 1), (1e9, 3, 5, 2)):
        o = ol.OracleAcqVariant(ol.ACQV_TONG, 4000, 4000, 5000, 250, 4000000, thr, tong_init_val=init, tong_max_val=mx, tong_max_dwells=20)
        o.set_local_code(code)
        o.restart()
        ev, rec = hl.acq_variant_attempt(service, 0, 3, code, 0, 4, sampled_ms=1, doppler_max=5000, doppler_step=250, fs_in=4000000, samples_per_ms=4000,
                                         samples_per_code=4000, threshold=thr, tong_init_val=init, tong_max_val=mx, tong_max_dwells=20)
        assert ev == want_event, (ev, rec)
        for k, row in enumerate(rec):
            ro = o.run(x4[4000 * k: 4000 * (k + 1)])
            assert int(row[1]) == ro.code_phase_index and row[2] == float(ro.doppler_hz) and row[7] == ro.acq_delay_samples, (k, row, ro.code_phase_index, ro.doppler_hz)
            assert close(row[3], ro.test_statistic) and close(row[4], ro.input_power), (k, row[3], ro.test_statistic)
            assert {0: 2, 1: 1, 2: 0}[int(row[0])] == ro.decision, (k, row[0], ro.decision)
        assert int(rec[-1][5]) == ro.tong_count and int(rec[-1][6]) == ro.dwell_count == len(rec)
        o.close()


def test_galileo_8ms_block_matches_oracle(service, golden_dir):
    """B200_Galileo_Pcps_8ms_Acquisition (codes A and B searched in ONE batch over shared forward spectra, 81 inclusive bins with
    the block's float32 phase-step rule) against the restated galileo_pcps_8ms_acquisition_cc, incl. the case where code B wins."""
    import gnss_b200 as gb
    import host_lib as hl

    xg = np.fromfile(golden_dir + "/Galileo_E1_ID_1_Fs_4Msps_8ms.dat", np.complex64)
    flipped = xg.copy()
    flipped[16000 + 2920:] *= -1  # data bit transition between the two code periods
    cg = synth.golden_array("gal_e1b_prn1_4msps_sampled")[:16000]
    code = np.concatenate([cg, cg])
    for stream_x, thr, want_event in ((xg, 1e-4, 1), (flipped, 1e-4, 1), (xg, 1e9, 2)):
        service.push(3, stream_x, gb.ITEM_32FC, 0)
        o = ol.OracleAcqVariant(ol.ACQV_8MS, 32000, 16000, 10000, 250, 4000000, thr, max_dwells=1)
        o.set_local_code(code)
        ro = o.run(stream_x)
        ev, rec = hl.acq_variant_attempt(service, 1, 3, code, 0, 1, sampled_ms=8, doppler_max=10000, doppler_step=250, fs_in=4000000, samples_per_ms=4000,
                                         samples_per_code=16000, threshold=thr, max_dwells=1)
        assert ev == want_event and len(rec) == 1, (ev, rec)
        row = rec[0]
        # The local replica is two identical code periods, so without a bit transition the correlation at lag 2920 and at lag
        # 2920 + 16000 is THE SAME NUMBER in exact arithmetic: which of the two a float32 transform reports first-and-largest is
        # rounding noise of that transform (the block's output, the lag modulo the code period, is the same either way).
        assert int(row[1]) % 16000 == ro.code_phase_index % 16000 == 2920 and row[2] == float(ro.doppler_hz) and row[7] == ro.acq_delay_samples, (row, ro.doppler_hz)
        if stream_x is flipped:
            assert int(row[1]) == ro.code_phase_index
        assert close(row[3], ro.test_statistic) and close(row[4], ro.input_power), (row[3], ro.test_statistic)
        o.close()


def test_e5a_noncoherent_iq_caf_block_matches_oracle(service):
    """B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf (up to four code spectra in ONE batch over shared forward spectra, per-bin
    A/B pick + I+Q row sum + row maximum on the device: gb200_acq_iq_combine) against the restated
    galileo_e5a_noncoherentIQ_acquisition_caf_cc: indices, bins, events and the Gnss_Synchro fields exactly, statistics to 1e-4,
    the CAF vector to 1e-4 of its largest magnitude."""
    import gnss_b200 as gb
    import host_lib as hl

    fs, spm, delay, dop = 16000000, 16000, 4321, 750.0
    x1 = synth.iq_pilot_scene(fs, spm, 5, 50.0, dop, delay, seed=11)
    x2 = synth.iq_pilot_scene(fs, 2 * spm, 5, 50.0, dop, delay, seed=12, data_flip_at=spm + delay, pilot_flip_at=spm + delay)
    noise = synth.make_iq(fs, spm, [], seed=13)
    c1 = synth.iq_pilot_codes(fs, spm, 5)
    c2 = synth.iq_pilot_codes(fs, 2 * spm, 5)
    c2z = synth.iq_pilot_codes(fs, 2 * spm, 5, zero_pad_from=spm)
    cases = [
        # (dwell streams, codes, sampled_ms, kwargs, threshold)
        ([x1], c1, 1, dict(), 1e-9),
        ([x1], c1, 1, dict(both_signal_components=True, caf_window_hz=1000), 1e-9),
        ([x1], c1, 1, dict(both_signal_components=True, caf_window_hz=1000), 1e9),
        ([x2], c2, 2, dict(caf_window_hz=1500), 1e-9),
        ([x2], c2, 2, dict(both_signal_components=True, caf_window_hz=1000), 1e-9),
        ([x2], c2z, 2, dict(both_signal_components=True, zero_padding=1), 1e-9),
        ([x1, noise], c1, 1, dict(bit_transition_flag=True, both_signal_components=True, max_dwells=2), 1e-9),
        ([noise, x1], c1, 1, dict(bit_transition_flag=True, max_dwells=2, caf_window_hz=1000), 1e-9),
    ]
    for streams, (ci, cq), sms, kw, thr in cases:
        okw = {k: v for k, v in kw.items() if k != "max_dwells"}
        o = ol.OracleE5aCaf(sms, spm, spm, kw.get("max_dwells", 1), 2000, 250, fs, thr, **okw)
        o.set_local_code(ci, cq)
        o.restart()
        want = [o.run(x) for x in streams]
        want_caf = o.caf_vector()
        service.push(3, np.concatenate(streams), gb.ITEM_32FC, 0)
        ev, rec, caf = hl.e5a_caf_attempt(service, 3, ci, cq, 0, len(streams), sampled_ms=sms, doppler_max=2000, doppler_step=250, fs_in=fs,
                                          samples_per_ms=spm, samples_per_code=spm, threshold=thr, **kw)
        assert len(rec) == len(want), (kw, rec)
        for row, ro in zip(rec, want):
            r = dict(zip(hl.E5A_RECORD_FIELDS, row))
            assert {0: 2, 1: 1, 2: 0}[int(r["event"])] == ro.decision, (kw, r, ro.decision)
            # with several code periods in the FFT the correlation is periodic in the code length (and a B replica peaks equally
            # at d and d + samples_per_code): the raw index is decided by rounding, the block itself reports it modulo the code
            assert int(r["code_phase_index"]) % spm == ro.code_phase_index % spm and int(r["doppler_bin"]) == ro.doppler_bin, (kw, r, ro.code_phase_index, ro.doppler_bin)
            assert sms > 1 or int(r["code_phase_index"]) == ro.code_phase_index
            assert r["doppler_hz"] == float(ro.doppler_hz) and r["acq_delay_samples"] == ro.acq_delay_samples, (kw, r, ro.doppler_hz)
            assert int(r["well_count"]) == ro.well_count and int(r["updated"]) == ro.updated, (kw, r)
            assert close(r["test_statistic"], ro.test_statistic) and close(r["input_power"], ro.input_power) and close(r["mag"], ro.mag), (kw, r, ro.mag)
        assert ev == {2: 0, 1: 1, 0: 2}[want[-1].decision]
        if kw.get("caf_window_hz", 0) > 0:
            assert np.max(np.abs(caf - want_caf)) <= 1e-4 * np.max(np.abs(want_caf)), (kw, caf, want_caf)
        o.close()
    # the first case really found the scene
    assert want is not None


def test_iq_combine_records_and_error_paths(service):
    """gb200_acq_iq_combine directly: per-bin records equal what the kept grids say (first maxima, A/B pick after the
    division by norm, the I_B-row read that stands in for magt_QB, maximum of the I+Q row), and it fails loudly (state error
    with a message) when a handle keeps no grid."""
    import gnss_b200 as gb

    fs, spm = 16000000, 16000
    x = synth.iq_pilot_scene(fs, 2 * spm, 9, 48.0, -500.0, 777, seed=21, data_flip_at=spm + 777)
    ci, cq = synth.iq_pilot_codes(fs, 2 * spm, 9)
    service.push(3, x, gb.ITEM_32FC, 0)
    steps = (2 * np.pi * np.arange(-1000, 1001, 250) / fs).astype(np.float32)
    codes = [ci, cq, np.concatenate([-ci[:spm], cq[spm:]]), np.concatenate([-cq[:spm], cq[spm:]])]
    hs = []
    for c in codes:
        a = gb.Acquisition(service, sampled_ms=2, ms_per_code=2, samples_per_chip=1, doppler_max=1000, doppler_step=250, fs_in=fs,
                           samples_per_ms=float(spm), samples_per_code=float(spm))
        a.set_doppler_grid(steps)
        a.set_local_code(c)
        hs.append(a)
    with pytest.raises(gb.Gb200Error) as e:
        gb.acq_iq_combine(*hs)
    assert e.value.code == -5 and "grid" in str(e.value)
    for a in hs:
        a.keep_grid(True)
    gb.acq_run_batch(hs, 3, 0)
    norm = np.float32(2 * spm) ** 4
    rec = gb.acq_iq_combine(*hs, norm=float(norm))
    grids = [a.read_grid() for a in hs]
    for b in range(len(steps)):
        rows = [g[b] for g in grids]
        idx = [int(np.argmax(r)) for r in rows]
        r = rec[b]
        assert list(r.index) == idx and [r.peak[k] for k in range(4)] == [rows[k][idx[k]] for k in range(4)]
        assert r.ib_at_qb == rows[2][idx[3]]
        m = [np.float32(rows[0][idx[0]]) / norm, np.float32(rows[1][idx[1]]) / norm, np.float32(rows[2][idx[2]]) / norm, np.float32(rows[2][idx[3]]) / norm]
        i_b, q_b = not (m[0] >= m[2]), not (m[1] >= m[3])
        assert r.picked == int(i_b) + 2 * int(q_b)
        comb = rows[2 if i_b else 0] + rows[3 if q_b else 1]
        assert r.indext == int(np.argmax(comb)) and r.mag == comb[r.indext]
    # the data-bit transition makes the B replica win on I in the signal's bin (-500 Hz -> bin 2)
    assert rec[2].picked & 1 and rec[2].indext % spm == 777
    for a in hs:
        a.close()


def test_fused_persistent_inverse_matches_default_path():
    """Tuning key acq_fused=1 (one persistent launch with a device-side tile queue instead of two kernels per chunk): a child
    process whose session service is steered by GB200_TUNE (tests/conftest.py -> gb200_service_set_tuning) runs the batch-search
    parity tests of this file under it; they compare with the oracle exactly as the default path does."""
    import subprocess
    import sys as _sys

    env = dict(os.environ, GB200_TUNE="acq_fused=1,acq_fused_lag=3,acq_fused_ring=5")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([_sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_acquisition.py"), "-m", "gpu", "-x", "-q", "-k",
                        "not fused and (fixture or batch or fast_inverse or ties or odd_lengths or zero_padded or bin_range)"], env=env, capture_output=True, text=True,
                       timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


@pytest.mark.parametrize("tune", ["acq_fwd_groups=4,acq_chunk_t=3", "acq_fwd_generic=1", "acq_variant=1"])
def test_forward_side_options_match_default_path(tune):
    """Tuning keys of the forward side, each in a child process steered by GB200_TUNE running this file's batch-search parity tests
    against the oracle: acq_fwd_groups > 1 (forward transforms in bin groups on their own stream, transforms numbered bin-major,
    inverse chunks waiting only for the groups they read; acq_chunk_t=3 forces several chunks on the small test grids) and
    acq_fwd_generic=1 (the run-time parametrised forward kernels where the compile-time DIF tiles would be used); acq_variant=1
    (the three-pass radix orders that were the defaults before the radix-20 / 25 butterflies, still selectable)."""
    import subprocess
    import sys as _sys

    env = dict(os.environ, GB200_TUNE=tune)
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([_sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_acquisition.py"), "-m", "gpu", "-x", "-q", "-k",
                        "not fused and not forward_side and (fixture or batch or fast_inverse or ties or odd_lengths or zero_padded or bin_range or multi_dwell)"], env=env,
                       capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


def test_real_capture_all_gps_prns(service, golden_dir):
    """The reference's real-signal fixture (src/tests/signal_samples/GSoC_CTTC_capture_2012_07_26_4Msps_4ms.dat: USRP1 roof capture;
    GPS PRN 29 and 25 stand out near +9.75 / +9 kHz, the front end's clock offset included): all 32 GPS PRNs searched in one
    batch on each of the four milliseconds, +-10 kHz / 250 Hz.  Against the oracle: input power and statistic to 1e-4 for every PRN and dwell; cell, bin and decision equal
    wherever the float32-DFT and float64-DFT oracles agree on the cell (a noise-only grid has near-ties no two FFTs order alike)."""
    import gnss_b200 as gb

    x = np.fromfile(golden_dir + "/GSoC_CTTC_capture_2012_07_26_4Msps_4ms.dat", np.complex64)
    assert len(x) == 16000
    service.push(2, x, gb.ITEM_32FC, 0)
    cfg = dict(doppler_max=10000, doppler_step=250, threshold=0.02, samples_per_chip=4)
    acqs = []
    for p in range(1, 33):
        a = gb.Acquisition(service, **cfg)
        a.set_local_code(gps_local_code(p))
        acqs.append(a)
    o32 = ol.OracleAcq(**cfg)
    o64 = ol.OracleAcq(fft_f64=1, **cfg)
    agreed, best = 0, (0.0, None)
    for ms in range(4):
        res = gb.acq_run_batch(acqs, 2, ms * 4000)
        seg = x[ms * 4000:(ms + 1) * 4000]
        for p, r in zip(range(1, 33), res):
            o32.set_local_code(gps_local_code(p))
            o64.set_local_code(gps_local_code(p))
            ra, rb = o32.run(seg, (ms + 1) * 4000), o64.run(seg, (ms + 1) * 4000)
            assert close(r.input_power, ra.input_power) and close(r.test_statistic, ra.test_statistic) and close(r.peak_mag, ra.peak_mag), (p, ms)
            if (ra.code_phase_index, ra.doppler_bin) == (rb.code_phase_index, rb.doppler_bin):
                agreed += 1
                check_against_oracle(r, ra, "prn %d ms %d" % (p, ms))
            if r.test_statistic > best[0]:
                best = (r.test_statistic, (p, r.doppler_hz, r.decision))
    assert agreed >= 100, agreed  # the oracles themselves agree on nearly every grid
    assert best[1][0] in (25, 29) and best[1][1] >= 9000 and best[1][2] == 1, best
    for a in acqs:
        a.close()
    o32.close()
    o64.close()


def test_service_coalesces_concurrent_acquisition_threads():
    """one thread per channel calling the blocking gb200_acq_run on the same sample block (what ChannelFsm starts
    together): with batch_window_us > 0 the service runs them as ONE gb200_acq_run_batch (forward spectra shared), every caller
    gets exactly the result of a lone call, and far fewer kernels are launched."""
    import threading

    import gnss_b200 as gb

    prns, x = c1_signal(8000)
    cfg = dict(doppler_max=5000, doppler_step=250, threshold=0.008, samples_per_chip=4)

    def run(window):
        svc = gb.Service(ring_capacity=1 << 20, ring_slack=1 << 16, batch_window_us=window)
        try:
            svc.push(0, x, gb.ITEM_32FC, 0)
            acqs = []
            for p in prns + [30, 31]:
                a = gb.Acquisition(svc, **cfg)
                a.set_local_code(gps_local_code(p))
                acqs.append(a)
            res = [None] * len(acqs)
            l0 = svc.launch_count()

            def worker(k):
                res[k] = acqs[k].run(0, 1000)

            th = [threading.Thread(target=worker, args=(k,)) for k in range(len(acqs))]
            for t in th:
                t.start()
            for t in th:
                t.join()
            launches = svc.launch_count() - l0
            stats = svc.acq_coalescing_stats()
            for a in acqs:
                a.close()
            return res, launches, stats
        finally:
            svc.close()

    alone, launches_alone, stats0 = run(0)
    assert stats0 == (0, 0)
    together, launches_together, (batches, requests) = run(20000)
    assert requests == len(alone) and batches < requests
    for ra, rt in zip(alone, together):
        for f in ("code_phase_index", "doppler_hz", "doppler_bin", "peak_mag", "input_power", "test_statistic", "decision", "acq_delay_samples",
                  "acq_samplestamp"):
            assert getattr(ra, f) == getattr(rt, f), f
    assert launches_together * 2 < launches_alone, (launches_together, launches_alone)
    assert sum(r.decision == 1 for r in together) == 8


def test_set_local_code_between_launch_and_collect(service):
    """The split launch / collect API keeps its results in a buffer of the batch's own: another channel's set_local_code (which
    reuses the service's staging area and waits for the stream) between the two calls must not disturb them."""
    import gnss_b200 as gb

    prns, x = c1_signal(8000, seed=314)
    service.push(2, x, gb.ITEM_32FC, 0)
    cfg = dict(doppler_max=5000, doppler_step=250, threshold=0.008, samples_per_chip=4)
    acqs = []
    for p in prns[:4]:
        a = gb.Acquisition(service, **cfg)
        a.set_local_code(gps_local_code(p))
        acqs.append(a)
    want = gb.acq_run_batch(acqs, 2, 100)
    other = gb.Acquisition(service, sampled_ms=4, ms_per_code=4, samples_per_ms=8000.0, samples_per_code=32000.0, fs_in=8000000, **cfg)
    handles = gb.acq_handles(acqs)
    gb.acq_launch_batch(acqs, handles, 2, 100)
    other.set_local_code(np.ones(32000, np.complex64))  # a larger staging request than the batch's: the old code reallocated here
    got = gb.acq_collect_batch(acqs, handles)
    for r, w in zip(got, want):
        assert (r.code_phase_index, r.doppler_bin, r.decision, r.peak_mag, r.test_statistic) == (w.code_phase_index, w.doppler_bin, w.decision, w.peak_mag, w.test_statistic)
    # a second launch of the same batch reuses the group tables on the device (same signature) and gives the same answer
    got2 = gb.acq_run_batch(acqs, 2, 100)
    for r, w in zip(got2, want):
        assert (r.code_phase_index, r.doppler_bin, r.decision, r.peak_mag) == (w.code_phase_index, w.doppler_bin, w.decision, w.peak_mag)
    # stop_acquisition(): launches abandoned without a wait leave no trace in the handles' state; a collected one after them is right
    for _ in range(3):
        gb.acq_launch_batch(acqs, handles, 2, 100)
        gb.acq_discard_batch(acqs, handles)
    with pytest.raises(gb.Gb200Error):
        gb.acq_discard_batch(acqs, handles)
    got2 = gb.acq_run_batch(acqs, 2, 100)
    for r, w in zip(got2, want):
        assert (r.code_phase_index, r.doppler_bin, r.decision, r.peak_mag) == (w.code_phase_index, w.doppler_bin, w.decision, w.peak_mag)
    # ... and after a Doppler-grid change of one member the tables are rebuilt
    acqs[1].set_doppler(2500, 250)
    got3 = gb.acq_run_batch(acqs, 2, 100)
    assert acqs[1].num_bins == 20 and got3[0].peak_mag == want[0].peak_mag and got3[1].doppler_bin < 20
    for a in acqs + [other]:
        a.close()


def test_two_threads_run_acquisitions_concurrently(service):
    """Two channel threads calling gb200_acq_run in a loop with no batching window: launch + collect of one call is one critical
    section, so neither ever sees the other's batch ('previous batch not collected') and both get their own results."""
    import threading

    import gnss_b200 as gb

    prns, x = c1_signal(8000, seed=2718)
    service.push(2, x, gb.ITEM_32FC, 0)
    cfg = dict(doppler_max=5000, doppler_step=250, threshold=0.008, samples_per_chip=4)
    acqs = []
    for p in (prns[0], prns[5]):
        a = gb.Acquisition(service, **cfg)
        a.set_local_code(gps_local_code(p))
        acqs.append(a)
    want = [a.run(2, 0) for a in acqs]
    errors = []

    def loop(k):
        try:
            for _ in range(40):
                r = acqs[k].run(2, 0)
                assert (r.code_phase_index, r.doppler_bin, r.peak_mag) == (want[k].code_phase_index, want[k].doppler_bin, want[k].peak_mag)
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=loop, args=(k,)) for k in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors[:3]
    for a in acqs:
        a.close()


def test_tuning_keys_and_stream_wait(service):
    import gnss_b200 as gb

    with pytest.raises(gb.Gb200Error):
        service.set_tuning("no_such_key", 1)
    with pytest.raises(gb.Gb200Error):
        service.set_tuning("acq_variant", 1)  # plans exist by now: must be set before the first acquisition handle
    service.set_tuning("acq_lanes", 1)
    service.set_tuning("acq_lanes", 2)
    x = np.zeros(1000, np.complex64)
    service.push(3, x, gb.ITEM_32FC, 0)
    service.wait_pushes(3)
    assert np.array_equal(service.read(3, 0, 1000), x)
