"""Host-side C++ mirror (gnss-sdr-1_b200/host/*.hpp) on CPU: the loop maths against the golden vectors generated from the
reference's own scalar libraries, and the DLL/PLL epoch loop (on the oracle correlator) locking onto a synthetic GPS L1 signal."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [HERE]
import host_lib as hl  # noqa: E402
import synth  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "loop_golden.npz"))


def test_discriminators_and_detectors_match_reference_golden(golden):
    L, p = hl.host(), golden["pts"]
    assert np.array_equal(np.array([L.gbh_pll_cloop_two_quadrant_atan(a[0], a[1]) for a in p]), golden["pll_two_quadrant"])
    assert np.array_equal(np.array([L.gbh_pll_four_quadrant_atan(a[0], a[1]) for a in p]), golden["pll_four_quadrant"])
    assert np.array_equal(np.array([L.gbh_fll_four_quadrant_atan(a[0], a[1], a[2], a[3], 0.0, 0.001) for a in p]), golden["fll_four_quadrant"])
    assert np.array_equal(np.array([L.gbh_dll_nc_e_minus_l_normalized(a[0], a[1], a[2], a[3]) for a in p]), golden["dll_nc_e_minus_l"])
    got = np.array([L.gbh_dll_nc_vemlp_normalized(a[0:2].copy(), a[2:4].copy(), a[4:6].copy(), a[6:8].copy()) for a in p])
    assert np.array_equal(got, golden["dll_nc_vemlp"])
    prompt = golden["prompt"].copy()
    assert np.float32(L.gbh_cn0_svn_estimator(prompt, 20, 0.001)) == golden["cn0"]
    assert np.float32(L.gbh_carrier_lock_detector(prompt, 20)) == golden["lock"]


def test_loop_filters_match_reference_golden(golden):
    L = hl.host()
    imp = np.zeros(8, np.float32)
    imp[1] = 1.0
    for order in (1, 2, 3):
        for last in (0, 1):
            y = np.zeros(8, np.float32)
            L.gbh_loop_filter_run(0.001, 5.0, order, last, 0.0, imp, y, 8)
            assert np.array_equal(y, golden["tlf_o%d_l%d" % (order, last)]), (order, last, y, golden["tlf_o%d_l%d" % (order, last)])
    for order in (2, 3):
        y = np.zeros(32, np.float32)
        L.gbh_fll_pll_filter_run(10.0, 35.0, order, 1234.5, golden["fll_disc"].copy(), golden["pll_disc"].copy(), 0.001, y, 32)
        assert np.array_equal(y, golden["fllpll_o%d" % order]), order


def gps_scene(n_ms, doppler, delay, cn0=47.0, prn=7, seed=5):
    fs = 4e6
    n = int(n_ms * 4000)
    x = synth.make_iq(fs, n, [dict(kind="gps_l1_ca", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay)], seed)
    return x, synth.code_table("gps_l1_ca", prn)


def test_closed_loop_locks_on_oracle_correlator():
    """start_tracking with an acquisition estimate 60 Hz and half a sample off; after pull-in the loop sits on the true Doppler,
    the code NCO on the true delay, C/N0 estimate near the simulated one, no loss of lock."""
    doppler, delay = 1830.0, 1517
    x, code = gps_scene(800, doppler, delay)
    conf = hl.gps_l1_conf()
    ep = hl.run_oracle_channel(x, 0, conf, code, acq_delay=float(delay) + 0.4, acq_doppler=doppler - 60.0, acq_stamp=0, n_steps=700)
    valid = [e for e in ep if e["valid"]]
    assert len(valid) > 600 and not any(e["loss_of_lock"] for e in ep)
    tail = valid[-200:]
    dop = np.array([e["carrier_doppler_hz"] for e in tail])
    assert abs(dop.mean() - doppler) < 2.0 and dop.std() < 6.0
    cn0 = np.array([e["cn0_db_hz"] for e in tail])
    assert abs(np.median(cn0) - 47.0) < 2.5
    # the code phase the loop reports: start of the PRN period in absolute samples, modulo one period
    starts = np.array([(e["tracking_sample_counter"] + e["code_phase_samples"]) % 4000.0 for e in tail])
    t = np.array([e["tracking_sample_counter"] for e in tail]) / 4e6
    true_start = delay - (doppler / 1575.42e6) * 4e6 * t  # the code Doppler of the simulated signal moves the period start
    err = (starts - true_start + 2000.0) % 4000.0 - 2000.0
    assert np.abs(err).max() < 0.25
    # prompt carries the power, early and late are balanced
    taps = np.array([e["taps"][:3] for e in tail])
    assert np.abs(taps[:, 1]).mean() > 1.6 * np.abs(taps[:, 0]).mean()
    assert abs(np.abs(taps[:, 0]).mean() - np.abs(taps[:, 2]).mean()) < 0.1 * np.abs(taps[:, 1]).mean()


# ---- states 3 / 4, secondary code, preamble, dump (8f N2) --------------------------------------------------

# log_data's record (dll_pll_veml_tracking.cc:1195-1243), the layout tracking_dump_reader.cc:38-60 reads back
DUMP_RECORD = np.dtype([("abs_VE", "<f4"), ("abs_E", "<f4"), ("abs_P", "<f4"), ("abs_L", "<f4"), ("abs_VL", "<f4"), ("prompt_I", "<f4"), ("prompt_Q", "<f4"),
                        ("PRN_start_sample_count", "<u8"), ("acc_carrier_phase_rad", "<f4"), ("carrier_doppler_hz", "<f4"),
                        ("carrier_doppler_rate_hz_s", "<f4"), ("code_freq_chips", "<f4"), ("code_freq_rate_chips", "<f4"), ("carr_error_hz", "<f4"),
                        ("carr_error_filt_hz", "<f4"), ("code_error_chips", "<f4"), ("code_error_filt_chips", "<f4"), ("CN0_SNV_dB_Hz", "<f4"),
                        ("carrier_lock_test", "<f4"), ("aux1", "<f4"), ("aux2", "<f8"), ("PRN", "<u4")])


def e1_scene(n_epochs, doppler, delay, cn0=46.0, fs=4e6, seed=20260002, prn=3):
    rng = np.random.default_rng(seed + 1)
    symbols = rng.integers(0, 2, 4096) * 2.0 - 1.0
    n = int(n_epochs * 0.004 * fs)
    x = synth.make_iq(fs, n, [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay,
                                   symbols=symbols, secondary=hl.GALILEO_E1_C_SECONDARY_CODE)], seed)
    return x, synth.code_table("gal_e1c_sinboc", prn), synth.code_table("gal_e1b_sinboc", prn), symbols


def test_e1_pilot_secondary_lock_extended_integration_and_dump(tmp_path):
    """Galileo E1 pilot tracking: wide tracking until the 25-chip secondary code is found in the prompt signs, then the 3 x (integrate)
    + 1 x (narrow DLL/PLL update) cycle with the narrow tap spacings; the data-prompt tap carries the E1B symbols; every DLL/PLL or
    integration cycle appends one 96-byte record to <dump_filename><channel>.dat."""
    doppler, delay, fs = -1270.0, 5333, 4e6
    x, pilot, data, symbols = e1_scene(430, doppler, delay)
    dump = str(tmp_path / "trk_dump_.dat")
    # 2nd-order PLL: the reference's 3rd-order carrier filter (tracking_FLL_PLL_filter.cc:105-118, mirrored bit-exactly) has a lightly damped
    # pole pair at (-0.1 +- 0.32j) w0 -- seconds of settling at these bandwidths, too long for a CPU test
    conf = hl.galileo_e1_conf(fs_in=fs, vector_length=16000, extend_correlation_symbols=4, pll_bw_hz=15.0, dll_bw_hz=2.0, pll_bw_narrow_hz=5.0,
                              dll_bw_narrow_hz=0.5, early_late_space_narrow_chips=0.1, very_early_late_space_narrow_chips=0.3, prn=3,
                              dump_filename=dump, cn0_min=25.0, pll_filter_order=2)
    ep = hl.run_oracle_channel(x, 0, conf, pilot, acq_delay=float(delay) + 0.3, acq_doppler=doppler + 20.0, acq_stamp=0, n_steps=420, data_code=data, channel=7)
    assert not any(e["loss_of_lock"] for e in ep)
    states = [e["state"] for e in ep]  # state AFTER the step
    first3 = states.index(3)
    # the window of 25 prompts has to fill (pull-in step + 25 epochs) and then align with the secondary code's start
    assert 25 <= first3 <= 130  # plus the PLL pull-in from the 20 Hz acquisition error
    assert set(states[:first3]) == {2}
    cyc = states[first3:]
    assert cyc[:9] == [3, 3, 3, 4, 3, 3, 3, 4, 3]  # (state after each step) 3 integrations in state 3, then the loop update in state 4
    assert all(e["correlation_length_ms"] == 4 for e in ep if e["valid"])
    # narrow spacing: the early/late taps sit closer to the prompt -> their magnitude ratio to the prompt rises from (1 - 3*0.15) to (1 - 3*0.1)
    # for a sinBOC(1,1) main peak (slope 3 per chip)
    wide = np.array([np.abs(e["taps"][:5]) for e in ep[first3 - 12:first3]])
    narrow = np.array([np.abs(e["taps"][:5]) for e in ep[-40:]])
    rw, rn = (wide[:, 1] + wide[:, 3]).mean() / (2 * wide[:, 2].mean()), (narrow[:, 1] + narrow[:, 3]).mean() / (2 * narrow[:, 2].mean())
    assert 0.45 < rw < 0.65 and rw + 0.1 < rn < 0.90  # (3.9 samples per chip: the sampled correlation function is a staircase)
    tail = [e for e in ep[-120:] if e["valid"]]
    dop = np.array([e["carrier_doppler_hz"] for e in tail])
    assert abs(dop.mean() - doppler) < 1.0 and dop.std() < 2.0
    # the estimate is refreshed every cn0_samples + 1 loop updates (84 epochs here); the last one is from narrow-tracking prompts only
    assert abs(tail[-1]["cn0_db_hz"] - 43.0) < 3.0  # the pilot carries half of the 46 dB-Hz
    # data-prompt tap: E1B symbols, up to the common sign of the carrier lock
    sc = np.array([e["tracking_sample_counter"] - e["current_prn_length_samples"] for e in tail], np.float64)
    period = np.round((sc - delay) / 16000.0).astype(int)
    got = np.sign([e["prompt_i"] for e in tail])
    want = symbols[period % len(symbols)]
    assert abs(np.mean(got * want)) > 0.98
    # dump: one record per valid epoch (state 2 cycles, integrations and narrow updates), 96 bytes each
    raw = np.fromfile(dump[:-4] + "7.dat", dtype=DUMP_RECORD)
    valid = [e for e in ep if e["valid"]]
    assert DUMP_RECORD.itemsize == 96 and len(raw) == len(valid)
    assert np.all(raw["PRN"] == 3)
    np.testing.assert_array_equal(raw["PRN_start_sample_count"], [e["tracking_sample_counter"] for e in valid])
    np.testing.assert_allclose(raw["carrier_doppler_hz"], [e["carrier_doppler_hz"] for e in valid], rtol=1e-6)
    np.testing.assert_allclose(raw["prompt_I"], [e["prompt_i"] for e in valid], rtol=1e-6)
    assert np.all(raw["aux2"] == raw["PRN_start_sample_count"].astype(np.float64))
    # narrow-update records hold the 4-epoch coherent sums: about 4x the single-epoch prompt magnitude
    st = np.array([e["state"] for e in valid])
    upd = raw["abs_P"][(st == 3) & (np.arange(len(st)) > first3 + 8)]
    single = np.abs([e["taps"][2] for e in valid[-40:]]).mean()
    assert 3.4 < np.median(upd) / single < 4.4


def test_gps_preamble_gates_extended_integration():
    """No secondary code (GPS L1 C/A): the loop stays in state 2 for 10 s, then waits for the telemetry preamble in the prompt signs
    (160 symbols) and only then starts the 20-symbol coherent integration, aligned with the bit edges."""
    fs, doppler, delay = 1.023e6, 740.0, 311
    bits = np.random.default_rng(5).integers(0, 2, 600) * 2 - 1
    for k in range(0, 600, 30):  # a preamble every 30 bits (0.6 s)
        bits[k:k + 8] = [1 if b else -1 for b in hl.GPS_PREAMBLE_BITS]
    n = int(11.9 * fs)
    x = synth.make_iq(fs, n, [dict(kind="gps_l1_ca", prn=9, cn0_dbhz=50.0, doppler_hz=doppler, delay_samples=delay, symbols=bits.astype(float),
                                   symbol_periods=20)], 99)
    conf = hl.gps_l1_conf(fs_in=fs, vector_length=1023, extend_correlation_symbols=20, pll_bw_hz=35.0, pll_bw_narrow_hz=10.0, dll_bw_narrow_hz=1.0,
                          early_late_space_narrow_chips=0.5, prn=9)
    ep = hl.run_oracle_channel(x, 0, conf, synth.code_table("gps_l1_ca", 9), acq_delay=float(delay), acq_doppler=doppler - 15.0, acq_stamp=0, n_steps=11800)
    assert not any(e["loss_of_lock"] for e in ep)
    states = np.array([e["state"] for e in ep])
    first3 = int(np.argmax(states == 3))
    t_switch = ep[first3]["tracking_sample_counter"] / fs
    assert 10.0 < t_switch < 10.0 + 0.6 + 0.17 and set(states[:first3]) == {2}
    # the preamble's last symbol is the last code period of a bit: the 20-symbol sums that follow never straddle a bit edge
    period = int(round((ep[first3]["tracking_sample_counter"] - ep[first3]["current_prn_length_samples"] - delay) / 1023.0))
    assert (period + 1) % 20 == 0
    upd = [i for i in range(first3 + 1, len(ep)) if states[i - 1] == 4 or (states[i] == 3 and states[i - 1] == 4)]
    four = [i for i in range(first3 + 1, len(ep)) if ep[i]["valid"] and states[i - 1] == 3 and states[i] == 4]
    assert len(four) > 50 and len(upd) > 0
    assert np.all(np.diff(four) == 20)
    dop = np.array([ep[i]["carrier_doppler_hz"] for i in four[-30:]])
    assert abs(dop.mean() - doppler) < 1.0


def test_high_dynamics_rate_smoothing_tracks_a_chirp():
    """high_dyn: the carrier phase-rate handed to the correlator comes from the two-window average of the NCO steps (:1018-1036);
    on a 400 Hz/s chirp it converges to 2 pi rate / fs^2 and the loop keeps lock."""
    fs, doppler, rate, delay = 4e6, 500.0, 400.0, 777
    n = int(1.2 * fs)
    x = synth.make_iq(fs, n, [dict(kind="gps_l1_ca", prn=4, cn0_dbhz=50.0, doppler_hz=doppler, delay_samples=delay, doppler_rate_hz_s=rate)], 41)
    conf = hl.gps_l1_conf(high_dyn=1, smoother_length=10, prn=4)
    ep = hl.run_oracle_channel(x, 0, conf, synth.code_table("gps_l1_ca", 4), acq_delay=float(delay), acq_doppler=doppler, acq_stamp=0, n_steps=1150)
    assert not any(e["loss_of_lock"] for e in ep)
    tail = [e for e in ep if e["valid"]][-300:]
    t = np.array([e["tracking_sample_counter"] for e in tail]) / fs
    dop = np.array([e["carrier_doppler_hz"] for e in tail])
    slope = np.polyfit(t, dop, 1)[0]
    assert abs(slope - rate) < 40.0
    assert np.abs(dop - (doppler + rate * t)).max() < 25.0


def test_threshold_from_pfa_is_the_exponential_quantile():
    """calculate_threshold (gps_l1_ca_pcps_acquisition.cc:262-279): inclusive bin count, ncells = vector_length * bins,
    threshold = quantile(Exp(lambda = vector_length), (1 - pfa)^(1 / ncells))."""
    from scipy.stats import expon

    L = hl.host()
    for pfa, dmax, step, n in [(0.001, 5000, 250, 4000), (0.01, 5000, 125, 32000), (1e-6, 10000, 500, 8000), (0.1, 5000, 250, 25000)]:
        bins = len(range(-dmax, dmax + 1, step))
        val = (1.0 - float(np.float32(pfa))) ** (1.0 / (n * bins))
        want = expon.ppf(val, scale=1.0 / n)
        got = L.gbh_pcps_threshold_from_pfa(pfa, dmax, step, n)
        assert got == pytest.approx(want, rel=1e-6)


def test_matfile_writer_is_read_back_by_scipy(tmp_path):
    """Level-5 MAT-file with the variable names / classes / dimensions of dump_results (pcps_acquisition.cc:486-535)."""
    from scipy.io import loadmat

    rng = np.random.default_rng(3)
    rows, cols = 1000, 7  # (effective_fft_size, num_doppler_bins): one COLUMN per Doppler bin
    grid = rng.random((cols, rows), dtype=np.float32)  # device layout [bin][index]
    fn = str(tmp_path / "acq_G_1C_ch_0_1_sat_9.mat")
    assert hl.host().gbh_matfile_write_grid(fn.encode(), grid.reshape(-1), rows, cols, 9, (1 << 40) + 5, 1) == 0
    m = loadmat(fn)
    assert m["acq_grid"].dtype == np.float32 and m["acq_grid"].shape == (rows, cols)
    np.testing.assert_array_equal(m["acq_grid"], grid.T)
    assert m["PRN"].dtype == np.uint32 and m["PRN"].shape == (1, 1) and int(m["PRN"][0, 0]) == 9
    assert m["sample_counter"].dtype == np.uint64 and int(m["sample_counter"][0, 0]) == (1 << 40) + 5
    assert m["d_positive_acq"].dtype == np.int32 and int(m["d_positive_acq"][0, 0]) == 1


def glonass_scene(n_ms, doppler, delay, freq_channel, cn0=47.0, fs=4e6, seed=20260007):
    x = synth.make_iq(fs, int(n_ms * fs / 1000), [dict(kind="glo_l1_ca", prn=1, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay,
                                                       carrier_offset_hz=562500.0 * freq_channel, phase=0.4)], seed)
    code = synth.code_table("glo_l1_ca", 1).astype(np.complex64)
    return x, code


def test_glonass_l1_loop_locks_on_oracle_correlator():
    """GlonassL1CaDllPllTracking (mirror of glonass_l1_ca_dll_pll_tracking_cc: FDMA carrier, complex code, 2nd-order DLL/PLL) on the
    oracle's Cpu_Multicorrelator restatement: pulls in from a coarse acquisition estimate and stays locked."""
    fs, doppler, delay, k = 4e6, 1234.0, 777.3, 1
    x, code = glonass_scene(700, doppler, delay, k)
    ep = hl.run_oracle_glonass(x, 0, code, k, acq_delay=delay + 0.6, acq_doppler=doppler - 40.0, acq_stamp=0, n_steps=680,
                               fs_in=fs, vector_length=4000, pll_bw_hz=40.0, dll_bw_hz=2.0, early_late_space_chips=0.5)
    F = {n: i for i, n in enumerate(hl.GLO_EPOCH_FIELDS)}
    assert ep[0, F["flags"]] == 0 and ep[0, F["prn_length_samples"]] == round(delay + 0.6 + 4000 - 0)  # pull-in alignment epoch
    tail = ep[-200:]
    assert np.all(tail[:, F["flags"]] == 1)  # valid symbols, no loss of lock
    assert abs(tail[:, F["carrier_doppler_hz"]].mean() - doppler) < 3.0
    assert 43.0 < tail[:, F["cn0_db_hz"]].mean() < 51.0
    assert np.all(tail[:, F["abs_p"]] > 1.5 * np.maximum(tail[:, F["abs_e"]], tail[:, F["abs_l"]]) * 0.5)
    # the block advances by whole code periods (4000 +- 1 samples at 4 Msps)
    assert set(np.unique(tail[:, F["prn_length_samples"]])) <= {3999.0, 4000.0, 4001.0}


def test_second_order_filters_match_reference_golden(golden_dir):
    """Tracking2ndDllFilter / Tracking2ndPllFilter (gb200_loop_math.hpp) bit-exact against the reference's
    Tracking_2nd_DLL_filter / Tracking_2nd_PLL_filter (tests/golden/loop2_golden.npz, generated from oracle/_ref)."""
    g = np.load(golden_dir + "/loop2_golden.npz")
    L = hl.host()
    for which, name in ((0, "dll"), (1, "pll")):
        for bw in (2.0, 50.0):
            y = np.zeros(64, np.float32)
            L.gbh_2nd_filter_run(which, bw, 0.001, g["disc"], y, 64)
            assert np.array_equal(y, g["%s_bw%g" % (name, bw)]), (name, bw)


def test_caf_filter_matches_oracle_arithmetic():
    """detail::caf_filter (host mirror of galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc:685-768) against the oracle's
    restatement, bit for bit: same integer promotions (the unsigned casts that make far-side weights hugely negative),
    same float32 accumulation order, with and without the Q component."""
    import oracle_lib as ol

    rng = np.random.default_rng(77)
    for bins, window, step in ((17, 1000, 250), (41, 3000, 250), (9, 500, 125), (81, 2000, 125)):
        vi = (rng.random(bins) * 1e9).astype(np.float32)
        vq = (rng.random(bins) * 1e9).astype(np.float32)
        for q in (None, vq):
            got = hl.caf_filter(vi, q, window, step)
            want = ol.e5a_caf_filter(vi, q, window, step)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (bins, window, step, q is None)


def test_acq_resampler_design_matches_oracle_restatement():
    """The flowgraph's acquisition resampler design (gnss_flowgraph.cc:431-447): the decimation rule and the firdes low-pass taps the
    C ABI produces against the oracle's restatement (bit for bit; GNU Radio itself is not in the tree, parity unpinned), plus
    what any correct design must show: odd symmetric taps, unit DC gain, > 45 dB rejection past cutoff + transition."""
    import gnss_b200 as gb
    import oracle_lib as ol

    L = ol.oracle()
    for fs, opt, want in ((4000000, 1000000, 4), (16000000, 1000000, 16), (25000000, 10000000, 2), (30000000, 10000000, 3), (6625000, 2000000, 2),
                          (12000000, 2000000, 6), (1000000, 1000000, 1), (8000000, 3000000, 2), (7000001, 2000000, 1), (500000, 1000000, 1)):
        assert hl.acq_resampler_decimation(fs, opt) == L.go_acq_resampler_decimation(fs, opt) == want, (fs, opt)
        assert fs % want == 0 and want <= max(1, fs // opt)
    for fs, dec in ((4e6, 4), (16e6, 4), (12e6, 6), (30e6, 3)):
        acq_fs = fs / dec
        got = gb.firdes_low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 10)
        want = ol.firdes_low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 10)
        assert len(got) == len(want) and len(got) % 2 == 1 and np.array_equal(got.view(np.uint32), want.view(np.uint32))
        assert len(got) == (int(53 * fs / (22 * acq_fs / 10)) | 1)
        assert np.array_equal(got, got[::-1]) and abs(float(got.astype(np.float64).sum()) - 1.0) < 1e-6
        h = np.abs(np.fft.rfft(got.astype(np.float64), 1 << 16))
        f = np.fft.rfftfreq(1 << 16, 1.0 / fs)
        assert 20 * np.log10(h[f > acq_fs / 2.1 + acq_fs / 10].max()) < -45.0
        assert abs(20 * np.log10(h[f < acq_fs / 2.1 - acq_fs / 10].min())) < 0.1
    # capacity protocol of the C entry point
    assert len(gb.firdes_low_pass(2.0, 4e6, 4e5, 2e5)) == (int(53 * 4e6 / (22 * 2e5)) | 1)


def caid_scene(n_ms, doppler, delay, cn0=48.0, fs=4e6, prn=9, seed=20260011):
    x = synth.make_iq(fs, int(n_ms * fs / 1000), [dict(kind="gps_l1_ca", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay, phase=0.7)], seed)
    return x, synth.code_table("gps_l1_ca", prn).astype(np.complex64)


@pytest.mark.parametrize("variant", [0, 1])
def test_c_aid_loop_locks_on_oracle_correlator(variant):
    """GpsL1CaDllPllCAidTracking (mirror of gps_l1_ca_dll_pll_c_aid_tracking_cc / _sc: FLL-assisted PLL filter with the FLL input at 0,
    2nd-order DLL, block length from the code NCO) on the oracle's Cpu_Multicorrelator / Cpu_Multicorrelator_16sc restatements:
    pull-in from a coarse estimate, lock, CN0; then extended integration once a preamble time is known (valid symbols every
    `extend_correlation_ms` blocks, NCOs free-running in between, lock kept)."""
    fs, doppler, delay = 4e6, -2222.0, 1500.4
    x, code = caid_scene(900, doppler, delay)
    # _sc: the int16 correlator accumulates in saturating int16 and the extended correlator adds prompts in int16, so the
    # front end has to deliver few-bit samples: noise sigma = 4 LSB keeps 5 coherent prompts (5 * 4000 * 0.18 * 4) inside +-32767
    samples = np.ascontiguousarray(np.rint(np.stack([x.real, x.imag], -1) * 4.0).astype(np.int16).reshape(-1)) if variant == 1 else x
    F = {n: i for i, n in enumerate(hl.CAID_EPOCH_FIELDS)}
    conf = dict(fs_in=fs, vector_length=4000, pll_bw_hz=35.0, dll_bw_hz=2.0, pll_bw_narrow_hz=15.0, dll_bw_narrow_hz=1.5, early_late_space_chips=0.5,
                extend_correlation_ms=5)
    ep = hl.run_oracle_caid(samples, variant, 0, code, acq_delay=delay + 0.7, acq_doppler=doppler + 60.0, acq_stamp=0, n_steps=850,
                            preamble=(500, 0.0), **conf)
    assert ep[0, F["flags"]] == 0 and ep[0, F["correlation_length_samples"]] == round(delay + 0.7 + 4000)  # pull-in alignment epoch
    lock = ep[300:500]
    assert np.all(lock[:, F["flags"]] == 1) and np.all(lock[:, F["correlation_length_ms"]] == 1)
    assert abs(lock[:, F["carrier_doppler_hz"]].mean() - doppler) < 4.0
    assert 44.0 < lock[:, F["cn0_db_hz"]].mean() < 52.0
    assert np.all(lock[:, F["abs_p"]] > 1.2 * np.maximum(lock[:, F["abs_e"]], lock[:, F["abs_l"]]))
    assert set(np.unique(lock[:, F["correlation_length_samples"]])) <= {3999.0, 4000.0, 4001.0}
    # extended integration: one valid symbol per 5 blocks with a 5x coherent prompt, Doppler still right
    ext = ep[600:]
    valid = ext[:, F["flags"]] == 1
    assert 0.18 < valid.mean() < 0.22 and np.all(ext[valid, F["correlation_length_ms"]] == 5)
    assert np.median(ext[valid, F["abs_p"]]) > 4.0 * np.median(lock[:, F["abs_p"]])
    assert abs(ext[valid, F["carrier_doppler_hz"]].mean() - doppler) < 4.0
    assert not np.any(ep[:, F["flags"]] >= 2)  # never lost lock
    if variant == 1:
        # the _sc correlator's outputs are integers
        assert np.all(ep[:, F["prompt_i"]] == np.round(ep[:, F["prompt_i"]]))
