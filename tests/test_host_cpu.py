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
