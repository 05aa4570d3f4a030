"""The reference's SIBLING blocks themselves as the oracle of their restatements and mirrors (CPU only).

pcps_tong_acquisition_cc, galileo_pcps_8ms_acquisition_cc, galileo_e5a_noncoherentIQ_acquisition_caf_cc, pcps_quicksync_acquisition_cc,
pcps_acquisition_fine_doppler_cc, glonass_l1_ca_dll_pll_tracking_cc, gps_l1_ca_dll_pll_c_aid_tracking_cc / _sc and
gps_l1_ca_kf_tracking_cc are compiled unmodified from /root/reference into oracle/_ref (oracle/ref_blocks2.cc drives them).
The restatements oracle/gnss_oracle.c go_acqv_* / go_e5a_* / go_qs_* / go_fd_* -- the checkers of the GPU tests of the B200_* acquisition
mirrors -- are held to the blocks dwell by dwell: statistic, input power, delay, Doppler, counters and events EQUAL (same DFT, same
float32 arithmetic); the tracker mirrors (host/gb200_*_tracking.hpp on the oracle's correlators) are held to the blocks epoch by
epoch.  This turns "block-level parity unpinned" into pinned for all of them."""
import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.skipif(ol.ref_blocks2() is None, reason="oracle/_ref (the compiled reference) not built")


def drive(block, x, vector_items, max_calls=200):
    """general_work() calls over `x` until an event is published; returns the calls that ran a dwell and the event"""
    pos, dwells = 0, []
    for _ in range(max_calls):
        n_avail = (len(x) - pos) // vector_items if vector_items > 1 else len(x) - pos
        r = block.work(x[pos:] if pos < len(x) else x[:1], max(0, n_avail))
        pos += r.consumed * (vector_items if vector_items > 1 else 1)
        dwells.append(r)
        if r.event:
            return dwells, r.event
    return dwells, 0


def test_tong_block_equals_restatement(golden_dir):
    x = np.fromfile(golden_dir + "/GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat", np.complex64)
    rng = np.random.default_rng(3)
    x4 = np.concatenate([x, x + 0.02 * (rng.standard_normal(8000) + 1j * rng.standard_normal(8000)).astype(np.complex64), x]).astype(np.complex64)
    code = synth.golden_array("gps_prn1_4msps_sampled")[:4000]
    for thr, init, mx, want_event in ((0.3, 1, 4, 1), (1e9, 3, 5, 2), (0.3, 2, 3, 1)):
        b = ol.RefAcqVariantBlock(ol.REF_TONG, 1, 5000, 250, 4000000, 4000, 4000, thr, tong_init_val=init, tong_max_val=mx, tong_max_dwells=20)
        b.set_local_code(code)
        b.set_state(1)
        o = ol.OracleAcqVariant(ol.ACQV_TONG, 4000, 4000, 5000, 250, 4000000, thr, tong_init_val=init, tong_max_val=mx, tong_max_dwells=20)
        o.set_local_code(code)
        o.restart()
        calls, ev = drive(b, x4, 4000)
        assert ev == want_event
        k = 0
        for r in calls:
            if r.state_before != 1:
                continue
            ro = o.run(x4[4000 * k: 4000 * (k + 1)])
            k += 1
            assert (r.test_statistics, r.input_power, r.acq_delay_samples, r.acq_doppler_hz, r.tong_count, r.count) == \
                (ro.test_statistic, ro.input_power, ro.acq_delay_samples, float(ro.doppler_hz), ro.tong_count, ro.dwell_count), (thr, k)
            assert {2: 1, 3: 0}.get(r.state_after, 2) == ro.decision
        assert k >= 1
        b.close()
        o.close()


def test_galileo_8ms_block_equals_restatement(golden_dir):
    xg = np.fromfile(golden_dir + "/Galileo_E1_ID_1_Fs_4Msps_8ms.dat", np.complex64)
    flipped = xg.copy()
    flipped[16000 + 2920:] *= -1  # data bit transition between the two code periods: code B wins
    cg = synth.golden_array("gal_e1b_prn1_4msps_sampled")[:16000]
    code = np.concatenate([cg, cg])
    for stream_x, thr, want_event in ((xg, 1e-4, 1), (flipped, 1e-4, 1), (xg, 1e9, 2)):
        b = ol.RefAcqVariantBlock(ol.REF_8MS, 8, 10000, 250, 4000000, 4000, 16000, thr, max_dwells=1, system=b"E", signal=b"1B")
        b.set_local_code(code)
        b.set_state(1)
        o = ol.OracleAcqVariant(ol.ACQV_8MS, 32000, 16000, 10000, 250, 4000000, thr, max_dwells=1)
        o.set_local_code(code)
        calls, ev = drive(b, np.concatenate([stream_x, stream_x]), 32000)
        ro = o.run(stream_x)
        assert ev == want_event == {1: 1, 0: 2}[ro.decision]
        r = [c for c in calls if c.state_before == 1][0]
        assert (r.test_statistics, r.input_power, r.acq_delay_samples, r.acq_doppler_hz) == (ro.test_statistic, ro.input_power, ro.acq_delay_samples, float(ro.doppler_hz))
        assert r.acq_delay_samples == 2920.0
        b.close()
        o.close()


E5A_FS, E5A_SPM = 16000000, 16000


@pytest.mark.parametrize("what", ["i_only", "i_and_q_caf", "two_periods_ab", "zero_padding", "bit_transition_two_dwells", "negative"])
def test_e5a_noncoherent_iq_caf_block_equals_restatement(what):
    delay, dop = 4321, 750.0
    kw = dict(bit_transition_flag=False, both_signal_components=False, caf_window_hz=0, zero_padding=0)
    sampled_ms, max_dwells, thr, n_dwells = 1, 1, 1e-9, 1
    x = synth.iq_pilot_scene(E5A_FS, E5A_SPM, 5, 50.0, dop, delay, seed=11)
    ci, cq = synth.iq_pilot_codes(E5A_FS, E5A_SPM, 5)
    if what == "i_and_q_caf":
        kw.update(both_signal_components=True, caf_window_hz=1000)
    elif what in ("two_periods_ab", "zero_padding"):
        sampled_ms = 2
        x = synth.iq_pilot_scene(E5A_FS, 2 * E5A_SPM, 5, 50.0, dop, delay, seed=12, data_flip_at=E5A_SPM + delay)
        ci, cq = synth.iq_pilot_codes(E5A_FS, 2 * E5A_SPM, 5, zero_pad_from=E5A_SPM if what == "zero_padding" else None)
        if what == "zero_padding":
            kw.update(zero_padding=1)
    elif what == "bit_transition_two_dwells":
        kw.update(bit_transition_flag=True)
        max_dwells, n_dwells = 2, 2
        rng = np.random.default_rng(5)
        x = np.concatenate([x, (rng.standard_normal(E5A_SPM) + 1j * rng.standard_normal(E5A_SPM)).astype(np.complex64)])
    elif what == "negative":
        thr = 1e9
    n = sampled_ms * E5A_SPM
    b = ol.RefAcqVariantBlock(ol.REF_E5A, sampled_ms, 2000, 250, E5A_FS, E5A_SPM, E5A_SPM, thr, max_dwells=max_dwells, system=b"E", signal=b"5X",
                              bit_transition_flag=int(kw["bit_transition_flag"]), both_signal_components=int(kw["both_signal_components"]),
                              caf_window_hz=kw["caf_window_hz"], zero_padding=kw["zero_padding"])
    b.set_local_code(ci, cq)
    b.set_state(1)
    o = ol.OracleE5aCaf(sampled_ms, E5A_SPM, E5A_SPM, max_dwells, 2000, 250, E5A_FS, thr, **kw)
    o.set_local_code(ci, cq)
    o.restart()
    calls, ev = drive(b, np.concatenate([x, x[:n]]), 1)
    cores = [c for c in calls if c.state_before == 2]
    assert len(cores) == n_dwells
    for k, r in enumerate(cores):
        ro = o.run(x[k * n: (k + 1) * n])
        assert (r.test_statistics, r.input_power, r.acq_delay_samples, r.acq_doppler_hz, r.count) == \
            (ro.test_statistic, ro.input_power, ro.acq_delay_samples, float(ro.doppler_hz), ro.well_count), (what, k)
    assert ev == {1: 1, 0: 2}[ro.decision]
    if kw["caf_window_hz"]:
        np.testing.assert_array_equal(b.caf_vector(o.num_bins), o.caf_vector())
    if what != "negative":
        assert ev == 1 and cores[0].acq_delay_samples % E5A_SPM == delay
    b.close()
    o.close()


# ------------------------------------------------------------------------------------------------ sibling trackers
def _same_rows(mirror, ref, len_field, F, what):
    import host_lib as hl  # noqa: F401

    n = min(len(mirror), len(ref))
    assert n > 0
    for i in range(n):
        a, r = mirror[i], ref[i]
        assert a[F[len_field]] == r[11], (what, i, "block length", a[F[len_field]], r[11])
        assert (int(a[F["flags"]]) & 1) == (int(r[10]) & 1), (what, i, "valid flag")
        if int(r[10]) & 1:
            for j in range(10):
                assert a[j] == r[j], (what, i, j, a[j], r[j])
    return n


def test_glonass_l1_tracker_mirror_equals_the_block():
    """GlonassL1CaDllPllTracking (host/gb200_glonass_tracking.hpp, on the oracle's complex-code correlator) against
    glonass_l1_ca_dll_pll_tracking_cc itself: FDMA channel +1 (PRN 1), 680 general_work calls, every field of every valid epoch and
    every block length equal."""
    import host_lib as hl
    import test_host_cpu as thc
    from test_ref_blocks_cpu import quiet

    fs, doppler, delay, k = 4e6, 1234.0, 777.3, 1
    x, code = thc.glonass_scene(700, doppler, delay, k)
    ep = hl.run_oracle_glonass(x, 0, code, k, acq_delay=delay + 0.6, acq_doppler=doppler - 40.0, acq_stamp=0, n_steps=680, fs_in=fs, vector_length=4000,
                               pll_bw_hz=40.0, dll_bw_hz=2.0, early_late_space_chips=0.5)
    with quiet():
        b = ol.RefSiblingTracker(ol.REF_TRK_GLONASS, fs, 4000, 40.0, 2.0, early_late_space_chips=0.5)
        b.start(1, delay + 0.6, doppler - 40.0, 0, 0)
        rr = b.run(x, 680)
        b.close()
    F = {n: i for i, n in enumerate(hl.GLO_EPOCH_FIELDS)}
    assert _same_rows(ep, rr, "prn_length_samples", F, "glonass") == 680 and (rr[:, 10].astype(int) & 1).sum() > 600


@pytest.mark.parametrize("variant", [0, 1])
def test_c_aid_tracker_mirrors_equal_the_blocks(variant):
    """GpsL1CaDllPllCAidTracking against gps_l1_ca_dll_pll_c_aid_tracking_cc (variant 0) / _sc (variant 1, int16 correlator and
    int16 extended-integration history): pull-in, 1 ms tracking, the preamble message after 500 blocks, then 5 ms extended
    integration -- 850 calls, all equal.  (This comparison found that only the _cc block retunes the DLL filter's integration
    time when it extends the correlation; the mirror had followed the _sc block for both.)"""
    import host_lib as hl
    import test_host_cpu as thc
    from test_ref_blocks_cpu import quiet

    fs, doppler, delay = 4e6, -2222.0, 1500.4
    x, code = thc.caid_scene(900, doppler, delay)
    samples = np.ascontiguousarray(np.rint(np.stack([x.real, x.imag], -1) * 4.0).astype(np.int16).reshape(-1)) if variant == 1 else x
    conf = dict(fs_in=fs, vector_length=4000, pll_bw_hz=35.0, dll_bw_hz=2.0, pll_bw_narrow_hz=15.0, dll_bw_narrow_hz=1.5, early_late_space_chips=0.5,
                extend_correlation_ms=5)
    ep = hl.run_oracle_caid(samples, variant, 0, code, acq_delay=delay + 0.7, acq_doppler=doppler + 60.0, acq_stamp=0, n_steps=850, preamble=(500, 0.0), **conf)
    with quiet():
        b = ol.RefSiblingTracker(ol.REF_TRK_CAID_CC + variant, fs, 4000, 35.0, 2.0, 15.0, 1.5, 0.5, 5)
        b.start(9, delay + 0.7, doppler + 60.0, 0, 0)
        rr = np.concatenate([b.run(samples, 500), (b.preamble(0.0), b.run(samples, 350))[1]])
        b.close()
    F = {n: i for i, n in enumerate(hl.CAID_EPOCH_FIELDS)}
    assert _same_rows(ep, rr, "correlation_length_samples", F, "c_aid %d" % variant) == 850
    valid = (rr[:, 10].astype(int) & 1) == 1
    assert valid[:500].sum() > 480 and 60 < valid[500:].sum() < 80 and np.all(rr[500:][valid[500:], 12] == 5)
    for i in np.nonzero(valid)[0]:
        assert ep[i, F["correlation_length_ms"]] == rr[i, 12]


@pytest.mark.parametrize("order,bce", [(2, False), (3, False), (2, True)])
def test_kf_tracker_mirror_equals_the_block(order, bce):
    """GpsL1CaKfTracking (host/gb200_kf_tracking.hpp, on the oracle's real-code correlator) against gps_l1_ca_kf_tracking_cc itself
    (compiled with the Armadillo stand-in of oracle/shim/armadillo, so the pin is on the block's own statement order, not on
    Armadillo's BLAS): 2-state and 3-state filters, and the 2-state filter with the Bayesian covariance estimator switched on after
    100 + 50 epochs -- pull-in epoch, 700 general_work calls, every field of every valid epoch and every block length equal."""
    import host_lib as hl
    import test_host_cpu as thc
    from test_ref_blocks_cpu import quiet

    fs, doppler, delay = 4e6, 1777.0, 2100.6
    x, code = thc.caid_scene(720, doppler, delay, cn0=47.0, prn=17, seed=20260031)
    kw = dict(order=order, bce_run=bce, bce_ptrans=100, bce_strans=50, bce_nu=10, bce_kappa=10)
    ep = hl.run_oracle_kf(x, 0, code.real, acq_delay=delay + 0.5, acq_doppler=doppler - 30.0, acq_stamp=0, acq_doppler_step=250, n_steps=700, fs_in=fs,
                          vector_length=4000, dll_bw_hz=2.0, early_late_space_chips=0.5, **kw)
    with quiet():
        b = ol.RefSiblingTracker(ol.REF_TRK_KF, fs, 4000, 0.0, 2.0, early_late_space_chips=0.5, **kw)
        b.start(17, delay + 0.5, doppler - 30.0, 0, 0, acq_doppler_step=250)
        rr = b.run(x, 700)
        b.close()
    F = {n: i for i, n in enumerate(hl.KF_EPOCH_FIELDS)}
    assert _same_rows(ep, rr, "prn_length_samples", F, "kf order %d bce %d" % (order, bce)) == 700
    lock = ep[400:]
    assert np.all(lock[:, F["flags"]] == 1) and abs(lock[:, F["carrier_doppler_hz"]].mean() - doppler) < 5.0
    assert 43.0 < lock[:, F["cn0_db_hz"]].mean() < 51.0
    if bce:  # the estimator replaced the CN0-derived measurement variance
        assert not np.array_equal(ep[300:, F["carr_phase_sigma2"]], hl.run_oracle_kf(
            x, 0, code.real, acq_delay=delay + 0.5, acq_doppler=doppler - 30.0, acq_stamp=0, acq_doppler_step=250, n_steps=700, fs_in=fs, vector_length=4000,
            dll_bw_hz=2.0, early_late_space_chips=0.5, order=order)[300:, F["carr_phase_sigma2"]])


# ------------------------------------------------------------------------------------------------ QuickSync
def _qs_scene(fs, spc, n_vectors, vec, delay, doppler, cn0, seed, prn=7):
    x = synth.make_iq(fs, n_vectors * vec, [dict(kind="gps_l1_ca", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay, phase=0.3)], seed)
    code1 = synth.sampled_code(synth.code_table("gps_l1_ca", prn), 1.023e6, fs, spc).astype(np.complex64)
    return x, code1


@pytest.mark.parametrize("what,folding,cn0,thr,max_dwells,bit_transition", [
    ("positive F=2", 2, 50.0, 0.05, 1, 0),
    ("positive F=4", 4, 52.0, 0.05, 1, 0),
    ("negative two dwells", 2, 0.0, 1e9, 2, 0),
    ("bit transition rule", 2, 50.0, 0.05, 2, 1),
])
def test_quicksync_block_equals_restatement(what, folding, cn0, thr, max_dwells, bit_transition):
    """pcps_quicksync_acquisition_cc itself against go_qs_*: folded search (folding_factor^2 segments onto samples_per_code /
    folding_factor points), input power over the used items, normalised maximum, the time-domain choice among the folding_factor
    candidate delays, well counter, sample stamp, state and event -- all EQUAL, dwell by dwell.  (The block adds its time-domain
    sums onto an uninitialised stack array; compiled here that memory is zero when the block reaches it, which is what the
    restatement assumes.)"""
    from test_ref_blocks_cpu import quiet

    fs, spc = 4_000_000, 4000
    sampled_ms = folding * 2  # adapter rule: a multiple of folding_factor (gps_l1_ca_pcps_quicksync_acquisition.cc:74-90)
    vec = sampled_ms * 4000
    delay, doppler = 2345, 1250.0
    x, code1 = _qs_scene(fs, spc, max_dwells + 2, vec, delay, doppler, cn0, 20260041 + folding)
    with quiet():
        b = ol.RefAcqVariantBlock(ol.REF_QUICKSYNC, sampled_ms, 5000, 250, fs, 4000, spc, thr, max_dwells=max_dwells, bit_transition_flag=bit_transition,
                                  folding_factor=folding, prn=7)
        code = np.tile(code1, sampled_ms // folding)  # the adapter's code_ (:243-247); the block copies samples_per_code items of it
        b.set_local_code(code)
        b.set_state(1)
        calls, ev = drive(b, x, vec)
    o = ol.OracleQuickSync(folding, sampled_ms, 5000, 250, fs, 4000, spc, thr, max_dwells=max_dwells, bit_transition_flag=bit_transition)
    o.set_local_code(code)
    o.restart()
    k = 0
    for r in calls:
        if r.state_before != 1:
            continue
        ro = o.run(x[k * vec: (k + 1) * vec], (k + 1) * vec)
        k += 1
        assert (r.test_statistics, r.input_power, r.mag, r.count, r.acq_delay_samples, r.acq_doppler_hz, r.acq_samplestamp_samples, r.state_after) == \
            (ro.test_statistic, ro.input_power, ro.mag, ro.well_count, ro.acq_delay_samples, ro.acq_doppler_hz, ro.acq_samplestamp, ro.state), (what, k)
    assert k == (max_dwells if ("negative" in what or bit_transition) else 1)
    assert ev == (2 if "negative" in what else 1)
    if "negative" not in what:
        assert ro.acq_delay_samples == delay and ro.acq_doppler_hz == doppler
    b.close()
    o.close()


# ------------------------------------------------------------------------------------------------ fine Doppler
@pytest.mark.parametrize("dop,delay,cn0,dwells,thr", [(1730.0, 1234, 50.0, 2, 2.5), (-3120.0, 3999, 47.0, 3, 2.5), (400.0, 0, 50.0, 1, 2.5), (0.0, 100, -100.0, 2, 2.5)])
def test_fine_doppler_block_equals_restatement(dop, delay, cn0, dwells, thr):
    """pcps_acquisition_fine_doppler_cc itself against go_fd_*: the un-normalised grid summed over max_dwells dwells, compute_CAF's
    first / second peak ratio, the 10 ms buffer, estimate_Doppler's replica (with its off-by-one std::rotate), the 8x zero-padded
    FFT's maximum, the bin-to-Hz table and the 1 kHz acceptance test -- statistic, delay, refined Doppler, sample stamp and event
    EQUAL (same scenes as tests/test_gpu_zz_fine_doppler.py, which holds the B200 mirror to go_fd_*)."""
    from test_ref_blocks_cpu import quiet

    fs, n = 4000000, 4000
    sats = [dict(kind="gps_l1_ca", prn=5, cn0_dbhz=cn0, doppler_hz=dop, delay_samples=delay)] if cn0 > 0 else []
    x = synth.make_iq(fs, 12 * n, sats, seed=int(abs(dop)) + 7)
    code = np.zeros(2 * n, np.float32)
    ol.oracle().go_gps_l1_ca_code_sampled(code, 5, fs, 0)
    code = code.view(np.complex64)
    o = ol.OracleFineDoppler(n, dwells, 5000, 500, fs, thr)
    ro = o.attempt(x, code)
    o.close()
    with quiet():
        b = ol.RefAcqVariantBlock(ol.REF_FINE_DOPPLER, 1, 5000, 500, fs, n, n, thr, max_dwells=dwells, prn=5)
        b.set_local_code(code)
        b.set_state(1)
        pos, ev, stat = 0, 0, None
        for _ in range(60):
            r = b.work(x[pos:], min(len(x) - pos, n))
            pos += r.consumed
            if r.state_before == 2:
                stat = r.test_statistics
            if r.event:
                ev = r.event
                break
        b.close()
    assert ev == (1 if ro.positive else 2) and stat == ro.test_statistic, (ev, stat, ro.test_statistic)
    if ro.positive:
        assert (r.acq_delay_samples, r.acq_doppler_hz) == (ro.acq_delay_samples, ro.acq_doppler_hz), (r.acq_delay_samples, r.acq_doppler_hz, ro.acq_doppler_hz)
        assert abs((r.acq_delay_samples - delay + n // 2) % n - n // 2) <= 1 and abs(r.acq_doppler_hz - dop) < 15.0 and ro.fine_accepted
