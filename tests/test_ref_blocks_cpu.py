"""The reference's GNU Radio blocks THEMSELVES as the oracle (CPU only, no GPU needed).

oracle/_ref/libgnss_ref.so now holds pcps_acquisition.cc and dll_pll_veml_tracking.cc compiled unmodified from /root/reference
against stand-in headers for the libraries that are not installed (oracle/shim: gr::block, pmt, a Boost subset, matio, armadillo,
upstream VOLK's generic kernels, and an FFT class on the oracle's DFT since FFTW is absent); oracle/ref_blocks.cc plays the GNU
Radio scheduler.  These tests pin, bit for bit,
  * the acquisition restatement (oracle/gnss_oracle.c go_acq_*, the checker of every GPU acquisition test) to the block:
    statistics, delay, Doppler, sample stamp, events and the whole magnitude grid over the block's feature matrix;
  * the host-side tracking mirror (gnss-sdr-1_b200/host/gb200_tracking.hpp, here on the oracle correlator) to the block:
    every field of every epoch through pull-in, wide tracking, secondary-code lock, extended integration and narrow tracking,
    loss of lock and the telemetry-fault message; and the 96-byte dump records, byte for byte, read back by the reference's own
    Tracking_Dump_Reader.
They are skipped only where oracle/_ref was never built (it is built in the container and travels to the GPU box)."""
import contextlib
import os

import numpy as np
import pytest

import host_lib as hl
import oracle_lib as ol
import synth

pytestmark = pytest.mark.skipif(ol.ref_blocks() is None, reason="oracle/_ref (the compiled reference) not built")


@contextlib.contextmanager
def quiet():
    """the reference blocks print 'Tracking of ... started' / 'Loss of lock' on std::cout"""
    fd = os.dup(1)
    null = os.open(os.devnull, os.O_WRONLY)
    try:
        os.dup2(null, 1)
        yield
    finally:
        os.dup2(fd, 1)
        os.close(null)
        os.close(fd)


# ------------------------------------------------------------------------------------------------ acquisition
def gps_code(prn, fs=4000000, n=4000):
    c = np.zeros(2 * n, np.float32)
    ol.oracle().go_gps_l1_ca_code_sampled(c, prn, fs, 0)
    return c.view(np.complex64)


def scene(n, seed, cn0=58.0):
    prns = [1, 3, 6, 9, 12, 17, 22, 28]
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=cn0 - k, doppler_hz=-4310 + 1237 * k, delay_samples=137 + 487 * k) for k, p in enumerate(prns)]
    return synth.make_iq(4e6, n, sats, seed)


def run_both(cfg, code, x, dwells, cshort=False, keep_grid=True, first=0):
    """the same dwells through the reference block and the restatement; returns [(block state, event, oracle result)], grids"""
    host = synth.quantize(x, "16ic") if cshort else x
    b = ol.RefAcqBlock(cshort=int(cshort), **cfg)
    o = ol.OracleAcq(cshort=int(cshort), keep_grid=int(keep_grid), **cfg)
    assert (b.fft_size, b.consumed, b.num_bins) == (o.fft_size, o.consumed, o.num_bins)
    b.set_local_code(code)
    o.set_local_code(code)
    b.start(first)
    out = []
    for d in range(dwells):
        off = d * b.consumed
        seg = host[2 * off: 2 * (off + b.consumed)] if cshort else host[off: off + b.consumed]
        st0 = b.state()
        if not st0.active and not st0.step_two:
            b.reset()  # the attempt has ended: ChannelFsm::start_acquisition -> reset() starts the next one
        st, ev = b.dwell(seg)
        ro = o.run(seg, first + off + b.consumed)
        out.append((st, ev, ro))
    b.close()
    o.close()
    return out


def same_dwell(st, ev, ro, what):
    want_event = {1: 1, 0: 2, 2: 0, 3: 0}[ro.decision]
    assert ev == want_event, (what, ev, ro.decision)
    assert st.test_statistics == ro.test_statistic, (what, st.test_statistics, ro.test_statistic)
    assert st.input_power == ro.input_power, (what, st.input_power, ro.input_power)
    assert (st.acq_delay_samples, st.acq_doppler_hz, st.acq_samplestamp_samples) == (ro.acq_delay_samples, ro.acq_doppler_hz, ro.acq_samplestamp), \
        (what, st.acq_delay_samples, ro.acq_delay_samples, st.acq_doppler_hz, ro.acq_doppler_hz, st.acq_samplestamp_samples, ro.acq_samplestamp)
    assert bool(st.step_two) == (ro.decision == 3), what


def test_acquisition_restatement_equals_the_block_on_the_reference_fixtures(golden_dir):
    x = np.fromfile(golden_dir + "/GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat", np.complex64)
    cfg = dict(doppler_max=5000, doppler_step=100, threshold=0.001, samples_per_chip=4)
    (st, ev, ro), = run_both(cfg, synth.golden_array("gps_prn1_4msps_sampled"), x, 1)
    same_dwell(st, ev, ro, "gps fixture")
    assert ev == 1 and (st.acq_delay_samples, st.acq_doppler_hz) == (524.0, 1700.0)  # gps_l1_ca_pcps_acquisition_test.cc:349-356
    x = np.fromfile(golden_dir + "/Galileo_E1_ID_1_Fs_4Msps_8ms.dat", np.complex64)
    cfg = dict(sampled_ms=4, ms_per_code=4, doppler_max=10000, doppler_step=250, threshold=1e-4, samples_per_ms=4000.0, samples_per_code=16000.0,
               samples_per_chip=4, system=b"E", signal=b"1B")
    ocfg = {k: v for k, v in cfg.items() if k not in ("system", "signal")}
    b = ol.RefAcqBlock(**cfg)
    o = ol.OracleAcq(**ocfg)
    code = synth.golden_array("gal_e1b_prn1_4msps_sampled")
    b.set_local_code(code)
    o.set_local_code(code)
    b.start(0)
    st, ev = b.dwell(x)
    ro = o.run(x[:16000], 16000)
    same_dwell(st, ev, ro, "galileo fixture")
    assert st.acq_delay_samples == 2920.0 and st.acq_doppler_hz == -750.0  # galileo_e1_pcps_ambiguous_acquisition_test.cc:293-294
    b.close()
    o.close()


@pytest.mark.parametrize("what", ["cfar", "cfar_16ic", "negative", "peak_stat", "multi_dwell", "two_step", "zero_padded", "bit_transition", "bit_transition_peak"])
def test_acquisition_restatement_equals_the_block_feature_matrix(what):
    """every branch of acquisition_core / the statistics / the decision logic the GPU tests rely on the restatement for"""
    x = scene(20000, seed=77)
    base = dict(doppler_max=5000, doppler_step=250, samples_per_chip=4)
    dwells, cshort, code = 1, False, gps_code(28)
    if what == "cfar":
        cfg = dict(threshold=0.008, **base)
    elif what == "cfar_16ic":
        cfg, cshort = dict(threshold=0.008, **base), True
    elif what == "negative":
        cfg, code = dict(threshold=0.008, **base), gps_code(31)
    elif what == "peak_stat":
        cfg, dwells = dict(use_CFAR_algorithm_flag=0, threshold=2.5, **base), 2
    elif what == "multi_dwell":
        cfg, dwells = dict(max_dwells=3, threshold=1e9, doppler_max=5000, doppler_step=500, samples_per_chip=4), 3
    elif what == "two_step":
        cfg, dwells = dict(make_2_steps=1, num_doppler_bins_step2=4, doppler_step2=125.0, threshold=0.008, doppler_max=5000, doppler_step=500, samples_per_chip=4), 2
    elif what == "zero_padded":
        cfg, code = dict(sampled_ms=2, ms_per_code=1, threshold=0.005, **base), np.tile(gps_code(1), 2)
    elif what == "bit_transition":
        cfg, code = dict(bit_transition_flag=1, threshold=0.001, **base), np.concatenate([gps_code(1), np.zeros(4000, np.complex64)])
    else:
        cfg, code = dict(bit_transition_flag=1, use_CFAR_algorithm_flag=0, threshold=1.5, **base), np.concatenate([gps_code(1), np.zeros(4000, np.complex64)])
    res = run_both(cfg, code, x, dwells, cshort=cshort)
    for d, (st, ev, ro) in enumerate(res):
        same_dwell(st, ev, ro, "%s dwell %d" % (what, d))
    events = [r[1] for r in res]
    if what == "multi_dwell":
        assert events == [0, 0, 2]
    if what == "two_step":
        assert events == [0, 1] and [r[2].decision for r in res] == [3, 1]
    if what in ("cfar", "cfar_16ic", "zero_padded", "bit_transition"):
        assert events == [1]
    if what == "negative":
        assert events == [2]


def test_acquisition_grid_and_channel_fsm_path():
    """The whole magnitude grid of an unfinished attempt (max_dwells = 2, first dwell) is bit-identical; with a ChannelFsm set the
    positive acquisition is a direct Event_valid_acquisition() call instead of event 1 (pcps_acquisition.cc:418-441)."""
    x = scene(8000, seed=5)
    cfg = dict(max_dwells=2, threshold=1e9, doppler_max=5000, doppler_step=500, samples_per_chip=4)
    b = ol.RefAcqBlock(**cfg)
    o = ol.OracleAcq(keep_grid=1, **cfg)
    b.set_local_code(gps_code(6))
    o.set_local_code(gps_code(6))
    b.start(0)
    st, ev = b.dwell(x[:4000])
    o.run(x[:4000], 4000)
    assert ev == 0 and st.dwell_counter == 1
    g, go = b.grid(), o.grid()
    assert g.shape == go.shape == (20, 4000) and np.array_equal(g, go)
    b.close()
    o.close()
    b = ol.RefAcqBlock(use_fsm=1, threshold=0.008, doppler_max=5000, doppler_step=250, samples_per_chip=4)
    b.set_local_code(gps_code(6))
    b.start(0)
    st, ev = b.dwell(x[:4000])
    assert ev == 0 and st.fsm_valid_acquisitions == 1
    b.close()


def test_glonass_fdma_offset_in_the_block():
    """is_fdma() (pcps_acquisition.cc:277-293): signal 1G adds DFRQ1_GLO * GLONASS_PRN[PRN] to every bin's frequency when the local
    code is set; the restatement's freq_offset_hz is that number.  PRN 1 -> channel +1, PRN 2 -> -4 (GLONASS_L1_L2_CA.h GLONASS_PRN)."""
    fs, n = 6625000, 6625
    code = synth.sampled_code(synth.code_table("glo_l1_ca", 1), 0.511e6, fs, n)
    for prn, k in ((1, 1), (2, -4)):
        x = synth.make_iq(fs, 2 * n, [dict(kind="glo_l1_ca", prn=1, cn0_dbhz=56, doppler_hz=1300.0, carrier_offset_hz=562500.0 * k, delay_samples=2222)], 40 + prn)
        cfg = dict(doppler_max=5000, doppler_step=250, threshold=0.003, samples_per_chip=13, fs_in=fs, samples_per_ms=6625.0, samples_per_code=6625.0)
        b = ol.RefAcqBlock(system=b"R", signal=b"1G", prn=prn, **cfg)
        b.set_local_code(code)
        assert b.state().old_freq == 562500 * k
        o = ol.OracleAcq(freq_offset_hz=562500 * k, **cfg)
        o.set_local_code(code)
        b.start(0)
        st, ev = b.dwell(x)
        ro = o.run(x[:n], n)
        same_dwell(st, ev, ro, "fdma k=%d" % k)
        assert abs(st.acq_delay_samples - 2222) <= 1
        b.close()
        o.close()


def test_input_power_accumulator_orders():
    """volk_32f_accumulator_s32f: the block with the shim's generic order equals the restatement's default; with VOLK's a_avx order
    (eight partial sums) it equals accumulator_lanes = 8 -- and the two differ by 4e-4 on 80000 8-bit items (tests/test_gpu_configs.py)."""
    fs, n = 20_000_000, 80000
    x = synth.make_iq(fs, n, [dict(kind="gal_e1b_sinboc", prn=11, cn0_dbhz=50, doppler_hz=-3100.0, delay_samples=901)], 20260004)
    q = synth.quantize(x, "8ic").astype(np.float32).view(np.complex64)
    code = synth.sampled_code(synth.code_table("gal_e1b_sinboc", 11), 2 * 1.023e6, fs, n)
    cfg = dict(sampled_ms=4, ms_per_code=4, samples_per_chip=20, doppler_max=500, doppler_step=250, fs_in=fs, samples_per_ms=20000.0, samples_per_code=80000.0,
               threshold=0.0015)
    powers = {}
    for lanes in (0, 8):
        ol.ref_blocks().gbref_set_accumulator_lanes(lanes)
        try:
            b = ol.RefAcqBlock(system=b"E", signal=b"1B", **cfg)
            b.set_local_code(code)
            b.start(0)
            st, ev = b.dwell(q)
        finally:
            ol.ref_blocks().gbref_set_accumulator_lanes(0)
        o = ol.OracleAcq(accumulator_lanes=lanes, **cfg)
        o.set_local_code(code)
        ro = o.run(q, n)
        same_dwell(st, ev, ro, "lanes %d" % lanes)
        powers[lanes] = st.input_power
        b.close()
        o.close()
    exact = float(np.sum(np.abs(q.astype(np.complex128)) ** 2) / n)
    assert abs(powers[8] - exact) < 1e-6 * exact and 3e-4 < (exact - powers[0]) / exact < 6e-4


# ------------------------------------------------------------------------------------------------ tracking
EPOCH_FIELDS = (("tracking_sample_counter", "tracking_sample_counter"), ("prompt_i", "prompt_i"), ("prompt_q", "prompt_q"),
                ("carrier_doppler_hz", "carrier_doppler_hz"), ("code_phase_samples", "code_phase_samples"), ("carrier_phase_rads", "carrier_phase_rads"),
                ("cn0_db_hz", "cn0_db_hz"), ("state", "state_after"), ("code_freq_chips", "code_freq_chips"),
                ("current_prn_length_samples", "n_samples"), ("correlation_length_ms", "correlation_length_ms"))


def assert_epochs_equal(mirror_epochs, ref_epochs, min_epochs):
    valid = [e for e in mirror_epochs if e["valid"]]
    n = min(len(valid), len(ref_epochs))
    assert n >= min_epochs, (len(valid), len(ref_epochs))
    for k in range(n):
        a, r = valid[k], ref_epochs[k]
        for fa, fr in EPOCH_FIELDS:
            assert a[fa] == getattr(r, fr), (k, fa, a[fa], getattr(r, fr))
        taps = np.array(r.corr[:], np.float32).view(np.complex64)
        nt = 5 if a["taps"][4] != 0 or taps[4] != 0 else 3
        assert np.array_equal(a["taps"][:nt], taps[:nt]), (k, a["taps"][:nt], taps[:nt])
    return n


def gps_scene(n_ms, doppler, delay, cn0=47.0, prn=7, seed=5, fs=4e6, **kw):
    n = int(n_ms * fs / 1000)
    x = synth.make_iq(fs, n, [dict(kind="gps_l1_ca", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay, **kw)], seed)
    return x, synth.code_table("gps_l1_ca", prn)


def test_c1_closed_loop_mirror_equals_the_tracking_block():
    """Config C1: GPS L1 C/A, 4 Msps, the adapter's defaults (PLL 50 Hz 3rd order, DLL 2 Hz 2nd order, E-L 0.5 chip), 2000
    epochs from an acquisition estimate 60 Hz and 0.4 samples off: run_dll_pll, update_tracking_vars (incl. the float-stored remnant
    carrier phase), the CN0 / lock detectors and the Gnss_Synchro fields of the mirror are bit-identical to the block's."""
    doppler, delay = 1830.0, 1517
    x, code = gps_scene(2100, doppler, delay)
    ep = hl.run_oracle_channel(x, 0, hl.gps_l1_conf(prn=7), code, acq_delay=delay + 0.4, acq_doppler=doppler - 60.0, acq_stamp=0, n_steps=2050)
    with quiet():
        b = ol.RefTrkBlock(pll_bw_hz=50.0, dll_bw_hz=2.0, pll_filter_order=3, dll_filter_order=2, early_late_space_chips=0.5, cn0_min=30, carrier_lock_th=0.80)
        b.start(7, delay + 0.4, doppler - 60.0, 0, 0)
        er, ev = b.run(x, 2050)
        b.close()
    assert ev == [] and assert_epochs_equal(ep, er, 2000) >= 2000
    assert abs(np.mean([e.carrier_doppler_hz for e in er[-200:]]) - doppler) < 2.0


def test_acquisition_to_tracking_hand_over_with_a_late_start():
    """start_tracking long after the acquisition stamp (the pull-in alignment of general_work state 1, :1568-1600: float fmod of the
    elapsed samples) and from a non-zero stream position."""
    doppler, delay = -2210.0, 3021
    x, code = gps_scene(400, doppler, delay, seed=9)
    first = 123457
    ep = hl.run_oracle_channel(x[first:], first, hl.gps_l1_conf(prn=7), code, acq_delay=delay - 0.3, acq_doppler=doppler + 35.0, acq_stamp=8000, n_steps=300)
    with quiet():
        b = ol.RefTrkBlock(cn0_min=30, carrier_lock_th=0.80)
        b.start(7, delay - 0.3, doppler + 35.0, 8000, first)
        er, ev = b.run(x[first:], 300)
        b.close()
    assert_epochs_equal(ep, er, 250)


def e1_scene(n_epochs, doppler, delay, cn0=46.0, fs=4e6, seed=20260002, prn=3):
    rng = np.random.default_rng(seed + 1)
    symbols = rng.integers(0, 2, 4096) * 2.0 - 1.0
    n = int(n_epochs * 0.004 * fs)
    x = synth.make_iq(fs, n, [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay,
                                   symbols=symbols, secondary=hl.GALILEO_E1_C_SECONDARY_CODE)], seed)
    return x, synth.code_table("gal_e1c_sinboc", prn), synth.code_table("gal_e1b_sinboc", prn)


def test_e1_pilot_extended_integration_mirror_equals_the_block_and_dumps_are_identical(tmp_path):
    """Galileo E1 pilot, extend_correlation_symbols = 4: wide tracking, secondary-code lock (acquire_secondary), the state 3 / 4 cycle
    with save_correlation_results' sign-by-secondary-chip sums and the narrow tap spacings -- every epoch equal; and the .dat dump
    the mirror writes is BYTE-IDENTICAL to the one the block writes (log_data, :1128-1250), both read by Tracking_Dump_Reader."""
    doppler, delay, fs = -1270.0, 5333, 4e6
    x, pilot, data = e1_scene(430, doppler, delay)
    d_m, d_r = str(tmp_path / "mirror_.dat"), str(tmp_path / "block_.dat")
    conf = hl.galileo_e1_conf(fs_in=fs, vector_length=16000, extend_correlation_symbols=4, pll_bw_hz=15.0, dll_bw_hz=2.0, pll_bw_narrow_hz=5.0,
                              dll_bw_narrow_hz=0.5, early_late_space_narrow_chips=0.1, very_early_late_space_narrow_chips=0.3, prn=3, cn0_min=25.0,
                              pll_filter_order=2, dump_filename=d_m)
    ep = hl.run_oracle_channel(x, 0, conf, pilot, acq_delay=delay + 0.3, acq_doppler=doppler + 20.0, acq_stamp=0, n_steps=420, data_code=data, channel=0)
    with quiet():
        b = ol.RefTrkBlock(fs_in=fs, vector_length=16000, system=b"E", signal=b"1B", pll_bw_hz=15.0, dll_bw_hz=2.0, pll_filter_order=2, dll_filter_order=2,
                           early_late_space_chips=0.15, very_early_late_space_chips=0.6, early_late_space_narrow_chips=0.1,
                           very_early_late_space_narrow_chips=0.3, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.5, extend_correlation_symbols=4, cn0_min=25,
                           carrier_lock_th=0.85, track_pilot=1, dump=1, dump_filename=d_r.encode())
        b.start(3, delay + 0.3, doppler + 20.0, 0, 0)
        er, ev = b.run(x, 420)
        b.close()
    n = assert_epochs_equal(ep, er, 410)
    assert {e.state_after for e in er} == {2, 3, 4}
    raw_m = open(d_m[:-4] + "0.dat", "rb").read()
    raw_r = open(d_r[:-4] + "0.dat", "rb").read()
    k = min(len(raw_m), len(raw_r)) // 96 * 96
    assert k >= 96 * (n - 1) and raw_m[:k] == raw_r[:k]
    nm, rec_m = ol.read_tracking_dump(d_m[:-4] + "0.dat")
    assert nm == len(raw_m) // 96 and np.array_equal(rec_m[:, 7], [e.tracking_sample_counter for e in er[: len(rec_m)]])
    np.testing.assert_allclose(rec_m[:, 9], [e.carrier_doppler_hz for e in er[: len(rec_m)]], rtol=1e-6)


def test_high_dynamics_and_fll_pull_in_mirror_equals_the_block():
    """high_dyn = true (phase / code rate smoothing over 2 x smoother_length epochs, the quadratic NCO terms, :1013-1063) with the FLL
    pull-in branch of run_dll_pll (:930-947) on a 300 Hz/s chirp."""
    doppler, delay = 900.0, 811
    x, code = gps_scene(1200, doppler, delay, cn0=50.0, prn=4, seed=11, doppler_rate_hz_s=300.0)
    conf = hl.gps_l1_conf(high_dyn=1, smoother_length=10, prn=4, enable_fll_pull_in=1, pull_in_time_s=1.0, fll_bw_hz=20.0)
    ep = hl.run_oracle_channel(x, 0, conf, code, acq_delay=float(delay), acq_doppler=doppler + 25.0, acq_stamp=0, n_steps=1150)
    with quiet():
        b = ol.RefTrkBlock(high_dyn=1, smoother_length=10, enable_fll_pull_in=1, pull_in_time_s=1, fll_bw_hz=20.0, cn0_min=30, carrier_lock_th=0.80)
        b.start(4, float(delay), doppler + 25.0, 0, 0)
        er, ev = b.run(x, 1150)
        b.close()
    assert_epochs_equal(ep, er, 1100)


def test_loss_of_lock_and_telemetry_fault():
    """Noise only: once the pull-in time has passed (an INTEGER number of elapsed seconds, :1555-1560) every failed lock test counts,
    max_lock_fail + 1 of them give event 3 and state 0 -- at the same epoch in the mirror; and the telemetry_to_trk fault message
    (int 1, :516-536) forces the same on a healthy channel at the next lock test."""
    fs = 4e6
    x = synth.make_iq(fs, int(1.7 * fs), [], 3)
    conf = hl.gps_l1_conf(prn=7, pull_in_time_s=0.0, max_lock_fail=10)
    ep = hl.run_oracle_channel(x, 0, conf, synth.code_table("gps_l1_ca", 7), acq_delay=100.0, acq_doppler=500.0, acq_stamp=0, n_steps=1650)
    with quiet():
        b = ol.RefTrkBlock(pull_in_time_s=0, max_lock_fail=10, cn0_min=30, carrier_lock_th=0.80)
        b.start(7, 100.0, 500.0, 0, 0)
        er, ev = b.run(x, 1650)
        st = b.state()
        b.close()
    assert ev == [3] and st == 0
    lost = [k for k, e in enumerate(ep) if e["loss_of_lock"]]
    assert len(lost) == 1 and ep[lost[0]]["state"] == 0
    assert_epochs_equal(ep[: lost[0]], er, 1000)
    assert len([e for e in ep[: lost[0]] if e["valid"]]) == len(er) and er[-1].tracking_sample_counter > fs
    # telemetry fault on a tracked satellite
    doppler, delay = 1830.0, 1517
    x, code = gps_scene(400, doppler, delay)
    ep = hl.run_oracle_channel(x, 0, hl.gps_l1_conf(prn=7), code, acq_delay=float(delay), acq_doppler=doppler, acq_stamp=0, n_steps=390, telemetry_fault_at=200)
    with quiet():
        b = ol.RefTrkBlock(cn0_min=30, carrier_lock_th=0.80)
        b.start(7, float(delay), doppler, 0, 0)
        er, ev = b.run(x, 199)  # the pull-in call produces no item: 199 items = 200 general_work calls
        b.telemetry_fault()
        er2, ev2 = b.run(x, 200)
        st = b.state()
        b.close()
    assert ev == [] and ev2 == [3] and st == 0
    lost = [k for k, e in enumerate(ep) if e["loss_of_lock"]]
    assert len(lost) == 1 and 200 <= lost[0] <= 221
    assert_epochs_equal(ep[: lost[0]], er + er2, 199)
    assert len([e for e in ep[: lost[0]] if e["valid"]]) == len(er) + len(er2)


@pytest.mark.parametrize("what", ["gps_l5i", "beidou_b1i", "galileo_e5a_i"])
def test_other_signals_mirror_equals_the_tracking_block(what):
    """The signals behind the table-driven adapters (gnss-sdr-1_b200/adapters/b200_signal_adapters.*): GPS L5 (data component, 10-bit
    Neuman-Hofman code: secondary-code lock, then the state 3 / 4 cycle over the 10 symbols of a bit), BeiDou B1I (20-bit NH code) and
    Galileo E5a-I (no secondary-code lock on the data component) with their adapters' default loop settings -- every epoch of the
    mirror equal to the block's.  (The block's own CN0 estimate on these 50 dB-Hz scenes is 33-40 dB-Hz once it sums prompts over a
    bit; the adapter tests on the GPU therefore do not compare CN0 with the scene's for these signals.)"""
    if what == "gps_l5i":
        fs, vlen, prn, sec = 12e6, 12000, 3, "0000110101"
        code = synth.code_table("gps_l5i", prn)
        sat = dict(kind="gps_l5i", prn=prn)
        sig = dict(signal_carrier_freq=1176.45e6, code_chip_rate=10.23e6, code_length_chips=10230, secondary=1, symbols_per_bit=10, secondary_code=sec)
        loop = dict(pll_bw_hz=50.0, dll_bw_hz=2.0, pll_bw_narrow_hz=2.0, dll_bw_narrow_hz=0.25, early_late_space_chips=0.5, early_late_space_narrow_chips=0.15,
                    carrier_lock_th=0.75)
        blk = dict(system=b"G", signal=b"L5")
    elif what == "beidou_b1i":
        fs, vlen, prn, sec = 8e6, 8000, 8, "00000100110101001110"
        code = synth.code_table("bds_b1i", prn)
        sat = dict(kind="bds_b1i", prn=prn)
        sig = dict(signal_carrier_freq=1561.098e6, code_chip_rate=2.046e6, code_length_chips=2046, secondary=1, symbols_per_bit=20, secondary_code=sec)
        loop = dict(pll_bw_hz=50.0, dll_bw_hz=2.0, pll_bw_narrow_hz=20.0, dll_bw_narrow_hz=2.0, early_late_space_chips=0.5, early_late_space_narrow_chips=0.5,
                    carrier_lock_th=0.85)
        blk = dict(system=b"C", signal=b"B1")
    else:
        R = ol.ref()
        fs, vlen, prn, sec = 12e6, 12000, 11, "10000100001011101001"
        c = np.zeros(2 * 10230, np.float32)
        R.gbref_galileo_e5a_code_primary(c, prn, b"5I")
        code = np.ascontiguousarray(c[0::2])
        sat = dict(kind="e5a_i", prn=prn, table=code, table_rate=10.23e6, carrier_hz=1176.45e6)
        sig = dict(signal_carrier_freq=1176.45e6, code_chip_rate=10.23e6, code_length_chips=10230, secondary=0, symbols_per_bit=20, secondary_code="")
        loop = dict(pll_bw_hz=20.0, dll_bw_hz=20.0, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=2.0, early_late_space_chips=0.5, early_late_space_narrow_chips=0.15,
                    carrier_lock_th=0.85)
        blk = dict(system=b"E", signal=b"5X")
    doppler, delay, n_ep = 1520.0, 2345, 260
    x = synth.make_iq(fs, int((n_ep + 12) * vlen), [dict(cn0_dbhz=50.0, doppler_hz=doppler, delay_samples=delay, secondary=sec, **sat)], 20260071)
    conf = hl.gps_l1_conf(fs_in=fs, vector_length=vlen, very_early_late_space_chips=0.0, very_early_late_space_narrow_chips=0.0, cn0_min=25.0, prn=prn,
                          preamble_symbols=[], **sig, **loop)
    ep = hl.run_oracle_channel(x, 0, conf, code, acq_delay=delay + 0.6, acq_doppler=doppler - 15.0, acq_stamp=0, n_steps=n_ep)
    with quiet():
        b = ol.RefTrkBlock(fs_in=fs, vector_length=vlen, very_early_late_space_chips=0.0, very_early_late_space_narrow_chips=0.0, cn0_min=25, **blk, **loop)
        b.start(prn, delay + 0.6, doppler - 15.0, 0, 0)
        er, ev = b.run(x, n_ep)
        b.close()
    assert ev == [] and assert_epochs_equal(ep, er, n_ep - 10) >= n_ep - 10
    assert abs(np.mean([e.carrier_doppler_hz for e in er[-60:]]) - doppler) < 6.0
    if what != "galileo_e5a_i":
        assert {e.state_after for e in er} >= {2, 3} or {e.state_after for e in er} >= {2, 4}  # secondary-code lock was reached
