"""Closed-loop parity of the host-side C++ adapter mirror on the B200 (SURVEY.md 8d, config C1 shape): the same DLL/PLL epoch
loop (gnss-sdr-1_b200/host/gb200_tracking.hpp) run on the B200 multicorrelator through the C ABI and on the oracle correlator,
from the same acquisition estimate, compared epoch by epoch; plus the acquisition block mirror."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [HERE]
import host_lib as hl  # noqa: E402
import synth  # noqa: E402

pytestmark = pytest.mark.gpu

PRNS = [1, 3, 6, 9, 12, 17, 22, 28]


def c1_scene(n_ms, seed=20260001, cn0=47.0):
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=cn0 - k, doppler_hz=-4375.0 + 1250.0 * k, delay_samples=137 + 487 * k) for k, p in enumerate(PRNS)]
    return sats, synth.make_iq(4e6, int(n_ms * 4000), sats, seed)


@pytest.mark.parametrize("item", ["32fc", "16ic"])
def test_c1_closed_loop_matches_oracle_loop(service, item):
    import gnss_b200 as gb

    sats, x = c1_scene(1000)
    it = {"32fc": gb.ITEM_32FC, "16ic": gb.ITEM_16IC}[item]
    if item == "16ic":
        q = synth.quantize(x, item)
        service.push(1, q, it, 0)
        x = (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)  # what the device converts to
    else:
        service.push(1, x, it, 0)
    conf = hl.gps_l1_conf()
    n_steps = 900
    for k in (0, 3, 7):
        s = sats[k]
        code = synth.code_table("gps_l1_ca", s["prn"])
        args = dict(conf=conf, code=code, acq_delay=float(s["delay_samples"]) + 0.3, acq_doppler=s["doppler_hz"] + 40.0, acq_stamp=0, n_steps=n_steps)
        ref = hl.run_oracle_channel(x, 0, **args)
        got = hl.run_b200_channel(service, 1, it, 0, **args)
        assert len(got) == len(ref) == n_steps
        gv = [e for e in got if e["valid"]]
        rv = [e for e in ref if e["valid"]]
        assert len(gv) == len(rv) > 850 and not any(e["loss_of_lock"] for e in got)
        # every epoch: correlator outputs within 1e-4 of the largest tap while the two loops are still on identical NCO state,
        # and the loop outputs stay together over the whole run (feedback keeps re-synchronising them)
        first = 40
        for g, r in list(zip(gv, rv))[:first]:
            assert g["tracking_sample_counter"] == r["tracking_sample_counter"]
            scale = np.abs(r["taps"][:3]).max()
            assert np.abs(g["taps"][:3] - r["taps"][:3]).max() <= 2e-4 * scale
        dop_g = np.array([e["carrier_doppler_hz"] for e in gv])
        dop_r = np.array([e["carrier_doppler_hz"] for e in rv])
        pos_g = np.array([e["tracking_sample_counter"] + e["code_phase_samples"] for e in gv])
        pos_r = np.array([e["tracking_sample_counter"] + e["code_phase_samples"] for e in rv])
        cn0_g = np.array([e["cn0_db_hz"] for e in gv])
        cn0_r = np.array([e["cn0_db_hz"] for e in rv])
        ph_g = np.array([e["carrier_phase_rads"] for e in gv])
        ph_r = np.array([e["carrier_phase_rads"] for e in rv])
        d_dop, d_pos, d_cn0, d_ph = np.abs(dop_g - dop_r), np.abs(pos_g - pos_r), np.abs(cn0_g - cn0_r), np.abs(ph_g - ph_r)
        print("closed loop %s prn %d: doppler max %.3g p95 %.3g Hz, position max %.3g samples, cn0 max %.3g dB, phase max %.3g rad" % (
            item, s["prn"], d_dop.max(), np.percentile(d_dop, 95), d_pos.max(), d_cn0.max(), d_ph.max()))
        # Closed loops amplify the 1e-5-level tap differences whenever floor(K_blk) lands on the other side of an integer in one
        # of the two runs (the next block then starts one sample apart until the DLL pulls both back): transient, bounded.
        # Yardstick: the oracle loop against itself with the input perturbed at float32 rounding level (1e-7 relative).
        rng = np.random.default_rng(k)
        xp = (x * (1.0 + 1e-7 * rng.standard_normal(len(x)))).astype(np.complex64)
        rp = [e for e in hl.run_oracle_channel(xp, 0, **args) if e["valid"]]
        s_dop = np.abs(np.array([e["carrier_doppler_hz"] for e in rp]) - dop_r)
        print("   oracle self-sensitivity: doppler max %.3g p95 %.3g Hz" % (s_dop.max(), np.percentile(s_dop, 95)))
        assert d_dop.max() < max(2.0, 4 * s_dop.max()) and np.percentile(d_dop, 95) < max(0.5, 4 * np.percentile(s_dop, 95))
        assert d_pos.max() < 0.05
        assert d_cn0.max() < 0.2
        assert d_ph.max() < 0.1
        # and the loop really tracks: Doppler on the truth, lock kept
        assert abs(dop_g[-200:].mean() - s["doppler_hz"]) < 3.0


def test_e1_pilot_extended_integration_matches_oracle_loop(service, tmp_path):
    """C2 / C5 shape: Galileo E1 pilot channel, VEML taps + data-prompt tap, 16-bit samples at 8 Msps; secondary-code lock, then
    4-symbol coherent integration with the narrow tap spacings (gb200_trk_set_taps).  Same state sequence as the loop on the oracle
    correlator, trajectories together, identical dump record stream layout."""
    import gnss_b200 as gb
    import test_host_cpu as thc

    fs, doppler, delay = 8e6, 2210.0, 9123
    x, pilot, data, symbols = thc.e1_scene(255, doppler, delay, cn0=47.0, fs=fs, seed=20260005, prn=11)  # fits the session ring (2^23 items)
    q = synth.quantize(x, "16ic")
    service.push(2, q, gb.ITEM_16IC, 0)
    x = (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)
    kw = dict(fs_in=fs, vector_length=32000, extend_correlation_symbols=4, pll_bw_hz=15.0, dll_bw_hz=2.0, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.5,
              early_late_space_narrow_chips=0.06, very_early_late_space_narrow_chips=0.25, prn=11, pll_filter_order=2)
    args = dict(code=pilot, acq_delay=float(delay) + 0.4, acq_doppler=doppler - 12.0, acq_stamp=0, n_steps=245, data_code=data)
    ref = hl.run_oracle_channel(x, 0, hl.galileo_e1_conf(dump_filename=str(tmp_path / "ref_.dat"), **kw), channel=3, **args)
    got = hl.run_b200_channel(service, 2, gb.ITEM_16IC, 0, hl.galileo_e1_conf(dump_filename=str(tmp_path / "b200_.dat"), **kw), channel=3, **args)
    assert len(got) == len(ref) == 245 and not any(e["loss_of_lock"] for e in got)
    st_g, st_r = [e["state"] for e in got], [e["state"] for e in ref]
    assert st_g == st_r and 3 in st_g and 4 in st_g
    first3 = st_g.index(3)
    for g, r in list(zip(got, ref))[1:30]:
        scale = np.abs(r["taps"]).max()
        assert g["tracking_sample_counter"] == r["tracking_sample_counter"]
        assert np.abs(g["taps"] - r["taps"]).max() <= 2e-4 * scale
    gv, rv = [e for e in got if e["valid"]], [e for e in ref if e["valid"]]
    dop_g, dop_r = np.array([e["carrier_doppler_hz"] for e in gv]), np.array([e["carrier_doppler_hz"] for e in rv])
    pos_g = np.array([e["tracking_sample_counter"] + e["code_phase_samples"] for e in gv])
    pos_r = np.array([e["tracking_sample_counter"] + e["code_phase_samples"] for e in rv])
    print("E1 closed loop: first state 3 at epoch %d, doppler diff max %.3g Hz, position diff max %.3g samples" % (
        first3, np.abs(dop_g - dop_r).max(), np.abs(pos_g - pos_r).max()))
    assert np.abs(dop_g - dop_r).max() < 1.0 and np.abs(pos_g - pos_r).max() < 0.05
    assert abs(dop_g[-60:].mean() - doppler) < 1.0
    # narrow taps took effect on the device: E and L move up the correlation peak
    wide = np.array([np.abs(e["taps"][:5]) for e in got[first3 - 10:first3]])
    narrow = np.array([np.abs(e["taps"][:5]) for e in got[-40:]])
    assert (narrow[:, 1] + narrow[:, 3]).mean() / narrow[:, 2].mean() > (wide[:, 1] + wide[:, 3]).mean() / wide[:, 2].mean() + 0.3
    # dump streams: same record count, integer fields identical, float fields together
    d_g = np.fromfile(str(tmp_path / "b200_3.dat"), dtype=thc.DUMP_RECORD)
    d_r = np.fromfile(str(tmp_path / "ref_3.dat"), dtype=thc.DUMP_RECORD)
    assert len(d_g) == len(d_r) == len(gv)
    np.testing.assert_array_equal(d_g["PRN_start_sample_count"], d_r["PRN_start_sample_count"])
    np.testing.assert_array_equal(d_g["PRN"], d_r["PRN"])
    # closed loop: the two runs share every input but not the last bits of each correlation, and the loop feeds those back;
    # the prompt magnitudes stay together to a few 1e-3 (typical 1e-5), the open-loop taps above to 2e-4
    rel = np.abs(d_g["abs_P"] - d_r["abs_P"]) / d_r["abs_P"]
    print("abs_P relative difference: median %.2g, 99th percentile %.2g, max %.2g" % (np.median(rel), np.percentile(rel, 99), rel.max()))
    assert np.median(rel) < 2e-4 and rel.max() < 5e-3


def test_acquisition_block_mirror(service):
    """B200_Pcps_Acquisition driven like Channel drives its adapter: set_channel, set_gnss_synchro, set_doppler_step, set_threshold,
    init, set_local_code, reset, then blocks until an event; positive for a present PRN, negative for an absent one."""
    import gnss_b200 as gb
    import oracle_lib as ol

    sats, x = c1_scene(8, cn0=58.0)
    service.push(2, x, gb.ITEM_32FC, 1000000)
    L = hl.host()

    def search(prn, threshold, max_dwells=1):
        cfg = gb.AcqCfg(sampled_ms=1, ms_per_code=1, samples_per_chip=4, max_dwells=max_dwells, doppler_max=5000, doppler_step=250,
                        num_doppler_bins_step2=4, doppler_step2=125.0, fs_in=4000000, samples_per_ms=4000.0, samples_per_code=4000.0,
                        bit_transition_flag=0, use_CFAR_algorithm_flag=1, make_2_steps=0, item_type=gb.ITEM_32FC, threshold=threshold, freq_offset_hz=0)
        code = np.zeros(8000, np.float32)
        ol.oracle().go_gps_l1_ca_code_sampled(code, prn, 4000000, 0)
        delay, dop, stamp, stat, mag = C.c_double(), C.c_double(), C.c_uint64(), C.c_float(), C.c_int()
        ev = L.gbh_acq_search(service.h, C.byref(cfg), 2, code, 4000, 1000000, 4, C.byref(delay), C.byref(dop), C.byref(stamp), C.byref(stat), C.byref(mag))
        return ev, delay.value, dop.value, stamp.value, stat.value

    # statistics first (threshold ~0 -> everything "positive"), then a threshold between the absent PRN's and the present ones'
    stats = {p: search(p, 1e-9)[4] for p in (sats[0]["prn"], sats[4]["prn"], 31)}
    th = float(np.sqrt(stats[31] * min(stats[sats[0]["prn"]], stats[sats[4]["prn"]])))
    assert stats[31] < th < min(stats[sats[0]["prn"]], stats[sats[4]["prn"]])
    for k in (0, 4):
        ev, delay, dop, stamp, stat = search(sats[k]["prn"], th)
        assert ev == 1 and stamp == 1000000 + 4000 and stat > th
        assert abs(delay - sats[k]["delay_samples"]) <= 1.0 and abs(dop - sats[k]["doppler_hz"]) <= 125.0
    # max_dwells > 1 switches the block to the first/second-peak statistic (pcps_acquisition.cc:152-159), a ratio of order 2 for
    # an absent PRN: below threshold 5 the block asks for the second dwell (no event), consumes the next d_consumed_samples,
    # then reports negative
    ev, _, _, stamp, stat = search(31, 5.0, max_dwells=2)
    assert ev == 2 and 1.0 < stat < 5.0 and stamp == 1000000 + 8000


def test_acquisition_dump_matfile(service, tmp_path):
    """dump=true on the searched channel: one .mat per finished search named like dump_results does (pcps_acquisition.cc:462-477), holding
    the magnitude grid (effective_fft_size x num_doppler_bins, one column per bin) and the scalars acquisition_dump_reader.cc reads;
    the grid equals the oracle's grid."""
    from scipy.io import loadmat

    import gnss_b200 as gb
    import oracle_lib as ol

    sats, x = c1_scene(8, cn0=58.0)
    service.push(2, x, gb.ITEM_32FC, 2000000)
    L = hl.host()
    prn = sats[2]["prn"]
    cfg = gb.AcqCfg(sampled_ms=1, ms_per_code=1, samples_per_chip=4, max_dwells=1, doppler_max=5000, doppler_step=250, num_doppler_bins_step2=4,
                    doppler_step2=125.0, fs_in=4000000, samples_per_ms=4000.0, samples_per_code=4000.0, bit_transition_flag=0,
                    use_CFAR_algorithm_flag=1, make_2_steps=0, item_type=gb.ITEM_32FC, threshold=0.001, freq_offset_hz=0)
    code = np.zeros(8000, np.float32)
    ol.oracle().go_gps_l1_ca_code_sampled(code, prn, 4000000, 0)
    delay, dop, stamp, stat, mag = C.c_double(), C.c_double(), C.c_uint64(), C.c_float(), C.c_int()
    name = C.create_string_buffer(512)
    ev = L.gbh_acq_search_dump(service.h, C.byref(cfg), 2, code, 4000, 2000000, 4, C.byref(delay), C.byref(dop), C.byref(stamp), C.byref(stat),
                               C.byref(mag), str(tmp_path / "acq_dump").encode(), 5, b"G", b"1C", prn, name, 512)
    assert ev == 1
    fn = name.value.decode()
    assert fn == str(tmp_path / "acq_dump") + "_G_1C_ch_5_1_sat_%d.mat" % prn
    m = loadmat(fn)
    assert m["acq_grid"].shape == (4000, 40) and m["acq_grid"].dtype == np.float32
    assert int(m["doppler_max"][0, 0]) == 5000 and int(m["doppler_step"][0, 0]) == 250 and int(m["PRN"][0, 0]) == prn
    assert int(m["d_positive_acq"][0, 0]) == 1 and int(m["num_dwells"][0, 0]) == 1 and int(m["sample_counter"][0, 0]) == 2004000
    assert float(m["acq_doppler_hz"][0, 0]) == dop.value and float(m["acq_delay_samples"][0, 0]) == np.float32(delay.value)
    assert float(m["test_statistic"][0, 0]) == stat.value and float(m["threshold"][0, 0]) == np.float32(0.001)
    # peak where the result says, grid equal to the oracle's
    col, row = np.unravel_index(np.argmax(m["acq_grid"].T), (40, 4000))
    assert -5000 + 250 * col == int(dop.value) and row == int(delay.value)
    o = ol.OracleAcq(doppler_max=5000, doppler_step=250, threshold=0.001, keep_grid=1)
    o.set_local_code(code.view(np.complex64)[:4000].copy())
    r = o.run(x[:4000], 2004000)
    want = o.grid()
    o.close()
    np.testing.assert_allclose(m["acq_grid"].T, want, rtol=2e-4, atol=1e-6 * want.max())
    assert float(m["input_power"][0, 0]) == pytest.approx(r.input_power, rel=1e-5)


def test_glonass_l1_loop_on_b200_matches_oracle_loop(service):
    """C4's fourth constellation: the GLONASS L1 C/A DLL/PLL block mirror (complex-code correlator, FDMA carrier k * 562.5 kHz,
    second-order filters) closed on the B200 correlator vs the same loop on the oracle correlator, 8-bit samples at 4 Msps."""
    import gnss_b200 as gb
    import test_host_cpu as thc

    fs, doppler, delay, k = 4e6, -2210.0, 1533.7, -3
    x, code = thc.glonass_scene(500, doppler, delay, k, cn0=48.0, seed=20260008)
    q = synth.quantize(x, "8ic")
    service.push(2, q, gb.ITEM_8IC, 0)
    xq = (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)
    kw = dict(fs_in=fs, vector_length=4000, pll_bw_hz=40.0, dll_bw_hz=2.0, early_late_space_chips=0.5)
    args = dict(code_c64=code, freq_channel=k, acq_delay=delay - 0.5, acq_doppler=doppler + 35.0, acq_stamp=0, n_steps=480)
    ref = hl.run_oracle_glonass(xq, 0, **args, **kw)
    got = hl.run_b200_glonass(service, 2, gb.ITEM_8IC, 0, **args, **kw)
    F = {n: i for i, n in enumerate(hl.GLO_EPOCH_FIELDS)}
    assert np.array_equal(got[:, F["tracking_sample_counter"]], ref[:, F["tracking_sample_counter"]])
    assert np.array_equal(got[:, F["flags"]], ref[:, F["flags"]]) and np.all(got[-200:, F["flags"]] == 1)
    rel = np.abs(got[1:, F["abs_p"]] - ref[1:, F["abs_p"]]) / ref[1:, F["abs_p"]].max()
    dd = np.abs(got[:, F["carrier_doppler_hz"]] - ref[:, F["carrier_doppler_hz"]])
    print("GLONASS closed loop: doppler diff max %.3g Hz, prompt magnitude diff max %.3g of peak, final doppler %.1f Hz" % (
        dd.max(), rel.max(), got[-1, F["carrier_doppler_hz"]]))
    assert dd.max() < 0.5 and rel.max() < 2e-3
    assert abs(got[-150:, F["carrier_doppler_hz"]].mean() - doppler) < 3.0
    assert np.abs(got[:, F["code_phase_samples"]] - ref[:, F["code_phase_samples"]]).max() < 0.01


def test_stream_decimator_matches_oracle_fir(service):
    """gb200_stream_set_decimator (the flowgraph's fir_filter_ccf(decimation, taps) acquisition resampler, run on the device behind
    every push): ragged pushes (shorter than the filter, not multiples of the decimation), a discontinuity that restarts the
    history, ring wrap-around, against the oracle's sequential float32 FIR; tolerance 2e-6 of sum|taps| * max|x|."""
    import gnss_b200 as gb
    import oracle_lib as ol

    rng = np.random.default_rng(31)
    taps = gb.firdes_low_pass(1.0, 4e6, 1e6 / 2.1, 1e5)
    x = (rng.standard_normal(60000) + 1j * rng.standard_normal(60000)).astype(np.complex64)
    service.set_decimator(0, 1, 4, taps)
    at = 0
    for n in (1000, 37, 3, 4099, 1, 20000, 34860):
        service.push(0, x[at: at + n], gb.ITEM_32FC, at)
        at += n
    assert at == len(x)
    want = ol.fir_decimate(x, taps, 4)
    got = service.read(1, 0, len(want))
    tol = 2e-6 * float(np.abs(taps).sum()) * float(np.abs(x).max())
    assert np.abs(got - want).max() <= tol, (np.abs(got - want).max(), tol)
    with pytest.raises(gb.Gb200Error):  # nothing beyond the newest complete output
        service.read(1, 0, len(want) + 1)
    # discontinuity: the filter history and the destination stream restart; item k of the destination sits at source item 4k
    seg = (rng.standard_normal(5000) + 1j * rng.standard_normal(5000)).astype(np.complex64)
    first = 100003
    service.push(0, seg, gb.ITEM_32FC, first)
    z = np.concatenate([np.zeros(3, np.complex64), seg])  # absolute index 100000 onwards, zeros before the segment
    want = ol.fir_decimate(z, taps, 4)[1:]                # k = 25001 onwards
    got = service.read(1, 25001, len(want))
    assert np.abs(got - want).max() <= tol
    with pytest.raises(gb.Gb200Error):
        service.read(1, 25000, 1)
    # other decimation / taps, long enough to wrap the 2^23-item ring
    service.set_decimator(0, 1, 0)
    t2 = rng.standard_normal(10).astype(np.float32)
    service.set_decimator(0, 1, 3, t2)
    big = (rng.standard_normal(9000000) + 1j * rng.standard_normal(9000000)).astype(np.complex64)
    first = 3000000
    for k in range(3):
        service.push(0, big[k * 3000000:(k + 1) * 3000000], gb.ITEM_32FC, first + k * 3000000)
    want = ol.fir_decimate(big, t2, 3)  # source starts at a multiple of 3: k0 = 1000000
    got = service.read(1, 1000000, len(want))
    assert np.abs(got - want).max() <= 2e-6 * float(np.abs(t2).sum()) * float(np.abs(big).max())
    service.set_decimator(0, 1, 0)
    # only gr_complex streams have the resampler (gps_l1_ca_pcps_acquisition.cc:90-94)
    service.set_decimator(0, 1, 4, taps)
    with pytest.raises(gb.Gb200Error) as e:
        service.push(0, np.zeros(64, np.int16), gb.ITEM_16IC, 0)
    assert e.value.code == -6
    service.set_decimator(0, 1, 0)


def test_acquisition_through_resampler_matches_oracle(service):
    """GNSS-SDR.use_acquisition_resampler end to end: 16 Msps GPS L1 C/A scene -> attach_acq_resampler (firdes taps, decimation by 4)
    -> B200_Pcps_Acquisition on the 4 Msps stream with the adapter's resampled Acq_Conf -> Gnss_Synchro scaled back by
    resampler_ratio and the filter latency (pcps_acquisition.cc:756-762), against the oracle's FIR + pcps_acquisition restatement."""
    import gnss_b200 as gb
    import oracle_lib as ol

    fs, dec = 16000000, 4
    assert hl.acq_resampler_decimation(fs, 4000000) == dec
    sats = [dict(kind="gps_l1_ca", prn=7, cn0_dbhz=50.0, doppler_hz=1300.0, delay_samples=5003.0)]
    x = synth.make_iq(fs, 40000, sats, seed=41)
    latency = hl.attach_acq_resampler(service, 0, 1, fs, dec)
    taps = ol.firdes_low_pass(1.0, fs, (fs / dec) / 2.1, (fs / dec) / 10)
    assert latency == (len(taps) - 1) // 2
    service.push(0, x, gb.ITEM_32FC, 0)
    xd = ol.fir_decimate(x, taps, dec)
    rfs = fs // dec
    code = np.zeros(2 * 4000, np.float32)
    ol.oracle().go_gps_l1_ca_code_sampled(code, 7, rfs, 0)
    o = ol.OracleAcq(sampled_ms=1, ms_per_code=1, samples_per_chip=4, doppler_max=5000, doppler_step=250, fs_in=rfs, samples_per_ms=4000.0,
                     samples_per_code=4000.0, threshold=1e-9)
    o.set_local_code(code.view(np.complex64))
    ro = o.run(xd[2000:6000], 6000)  # the block's sample counter stands behind the consumed block
    cfg = gb.AcqCfg(sampled_ms=1, ms_per_code=1, samples_per_chip=4, max_dwells=1, doppler_max=5000, doppler_step=250, num_doppler_bins_step2=4,
                    doppler_step2=125.0, fs_in=rfs, samples_per_ms=4000.0, samples_per_code=4000.0, bit_transition_flag=0, use_CFAR_algorithm_flag=1,
                    make_2_steps=0, item_type=gb.ITEM_32FC, threshold=1e-9, freq_offset_hz=0)
    ev, delay, dop, stamp, stat = hl.acq_search_resampled(service, cfg, 1, code.view(np.complex64), 2000, 1, float(dec), latency)
    assert ev == 1 and dop == float(ro.doppler_hz) == 1250.0
    want_delay = float(np.fmod(np.float32(ro.code_phase_index), np.float32(4000.0))) * dec - latency
    assert delay == want_delay and stamp == int(np.rint(float(ro.acq_samplestamp) * dec))
    assert abs(stat - ro.test_statistic) <= 1e-4 * ro.test_statistic
    # and it is the right answer: the peak sits where the scene put the code (filter delay removed, +-2 resampled samples)
    assert abs((delay - (5003.0 - 2000 * dec)) % 16000) <= 2 * dec or abs((delay - (5003.0 - 2000 * dec)) % 16000 - 16000) <= 2 * dec
    service.set_decimator(0, 1, 0)
    o.close()


@pytest.mark.parametrize("variant", [0, 1])
def test_c_aid_loops_on_b200_match_oracle_loops(service, variant):
    """The carrier-aided GPS L1 C/A trackers (gps_l1_ca_dll_pll_c_aid_tracking_cc on the complex-code correlator, _sc on the int16
    correlator) closed on the B200 through the C ABI vs the same loop on the oracle correlators, incl. the switch to 5 ms
    extended integration.  The int16 correlator is bit-exact, so the whole _sc loop is: every field of every epoch equal."""
    import gnss_b200 as gb
    import test_host_cpu as thc

    fs, doppler, delay = 4e6, 3100.0, 2750.6
    x, code = thc.caid_scene(700, doppler, delay, seed=20260012)
    conf = dict(fs_in=fs, vector_length=4000, pll_bw_hz=35.0, dll_bw_hz=2.0, pll_bw_narrow_hz=15.0, dll_bw_narrow_hz=1.5, early_late_space_chips=0.5,
                extend_correlation_ms=5)
    args = dict(code_c64=code, acq_delay=delay - 0.6, acq_doppler=doppler - 50.0, acq_stamp=0, n_steps=650, preamble=(400, 0.0))
    F = {n: i for i, n in enumerate(hl.CAID_EPOCH_FIELDS)}
    if variant == 1:
        q = np.ascontiguousarray(np.rint(np.stack([x.real, x.imag], -1) * 4.0).astype(np.int16).reshape(-1))
        service.push(2, q, gb.ITEM_16IC, 0)
        ref = hl.run_oracle_caid(q, 1, 0, **args, **conf)
        got = hl.run_b200_caid(service, 2, 1, 0, **args, **conf)
        assert np.array_equal(got, ref)
    else:
        service.push(2, x, gb.ITEM_32FC, 0)
        ref = hl.run_oracle_caid(x, 0, 0, **args, **conf)
        got = hl.run_b200_caid(service, 2, 0, 0, **args, **conf)
        assert np.array_equal(got[:, F["tracking_sample_counter"]], ref[:, F["tracking_sample_counter"]])
        assert np.array_equal(got[:, F["flags"]], ref[:, F["flags"]])
        rel = np.abs(got[1:, F["abs_p"]] - ref[1:, F["abs_p"]]) / ref[1:, F["abs_p"]].max()
        dd = np.abs(got[:, F["carrier_doppler_hz"]] - ref[:, F["carrier_doppler_hz"]])
        print("c_aid _cc closed loop: doppler diff max %.3g Hz, prompt magnitude diff max %.3g of peak" % (dd.max(), rel.max()))
        assert dd.max() < 0.5 and rel.max() < 2e-3
        assert np.abs(got[:, F["code_phase_samples"]] - ref[:, F["code_phase_samples"]]).max() < 0.01
    lock = got[250:400]
    assert np.all(lock[:, F["flags"]] == 1) and abs(lock[:, F["carrier_doppler_hz"]].mean() - doppler) < 4.0
    ext = got[450:]
    assert np.all(ext[ext[:, F["flags"]] == 1, F["correlation_length_ms"]] == 5) and 0.18 < (ext[:, F["flags"]] == 1).mean() < 0.22


def test_c1_b200_loop_against_the_reference_tracking_block(service):
    """The same comparison with the reference's OWN block on the other side: dll_pll_veml_tracking.cc compiled from /root/reference
    (oracle/_ref, driven by oracle/ref_blocks.cc) against the mirror loop on the B200 correlator, C1 scene, 900 epochs: identical
    sample counters / block lengths while the NCO states coincide, correlator taps within 2e-4 of the largest tap, Doppler / code
    position / CN0 / carrier phase trajectories together."""
    import gnss_b200 as gb
    import oracle_lib as ol
    import test_ref_blocks_cpu as trb

    if ol.ref_blocks() is None:
        pytest.skip("oracle/_ref (the compiled reference) not built")
    sats, x = c1_scene(1000)
    service.push(1, x, gb.ITEM_32FC, 0)
    s = sats[2]
    code = synth.code_table("gps_l1_ca", s["prn"])
    acq_delay, acq_doppler = float(s["delay_samples"]) - 0.2, s["doppler_hz"] - 30.0
    got = hl.run_b200_channel(service, 1, gb.ITEM_32FC, 0, conf=hl.gps_l1_conf(prn=s["prn"]), code=code, acq_delay=acq_delay, acq_doppler=acq_doppler, acq_stamp=0,
                              n_steps=900)
    with trb.quiet():
        b = ol.RefTrkBlock(pll_bw_hz=50.0, dll_bw_hz=2.0, pll_filter_order=3, dll_filter_order=2, early_late_space_chips=0.5, cn0_min=30, carrier_lock_th=0.80)
        b.start(s["prn"], acq_delay, acq_doppler, 0, 0)
        er, ev = b.run(x, 900)
        b.close()
    gv = [e for e in got if e["valid"]]
    assert ev == [] and not any(e["loss_of_lock"] for e in got) and len(gv) >= 890 and len(er) >= len(gv)
    for g, r in list(zip(gv, er))[:40]:
        assert g["tracking_sample_counter"] == r.tracking_sample_counter and g["current_prn_length_samples"] == r.n_samples
        taps = np.array(r.corr[:6], np.float32).view(np.complex64)
        assert np.abs(g["taps"][:3] - taps).max() <= 2e-4 * np.abs(taps).max()
    n = len(gv)
    d_dop = np.abs(np.array([e["carrier_doppler_hz"] for e in gv]) - np.array([e.carrier_doppler_hz for e in er[:n]]))
    d_pos = np.abs(np.array([e["tracking_sample_counter"] + e["code_phase_samples"] for e in gv]) - np.array([e.tracking_sample_counter + e.code_phase_samples for e in er[:n]]))
    d_cn0 = np.abs(np.array([e["cn0_db_hz"] for e in gv]) - np.array([e.cn0_db_hz for e in er[:n]]))
    d_ph = np.abs(np.array([e["carrier_phase_rads"] for e in gv]) - np.array([e.carrier_phase_rads for e in er[:n]]))
    print("vs reference block: doppler max %.3g Hz, position max %.3g samples, cn0 max %.3g dB, phase max %.3g rad" % (d_dop.max(), d_pos.max(), d_cn0.max(), d_ph.max()))
    assert d_dop.max() < 2.0 and np.percentile(d_dop, 95) < 0.5 and d_pos.max() < 0.05 and d_cn0.max() < 0.2 and d_ph.max() < 0.1
    assert abs(np.mean([e["carrier_doppler_hz"] for e in gv[-200:]]) - s["doppler_hz"]) < 3.0


def test_acquisition_against_the_reference_block(service):
    """C1 batch against pcps_acquisition.cc itself (oracle/_ref): 8 present PRNs + 1 absent, gr_complex and cshort input."""
    import gnss_b200 as gb
    import oracle_lib as ol

    if ol.ref_blocks() is None:
        pytest.skip("oracle/_ref (the compiled reference) not built")
    sats, x = c1_scene(3, cn0=58.0)
    for item, it in (("32fc", gb.ITEM_32FC), ("16ic", gb.ITEM_16IC)):
        host = x if item == "32fc" else synth.quantize(x, "16ic")
        service.push(2, host, it, 777)
        cfg = dict(doppler_max=5000, doppler_step=250, threshold=0.008, samples_per_chip=4)
        acqs, codes = [], []
        for p in PRNS + [31]:
            c = np.zeros(8000, np.float32)
            ol.oracle().go_gps_l1_ca_code_sampled(c, p, 4000000, 0)
            codes.append(c.view(np.complex64))
            a = gb.Acquisition(service, item_type=it, **cfg)
            a.set_local_code(codes[-1])
            acqs.append(a)
        res = gb.acq_run_batch(acqs, 2, 777 + 500)
        for k, (a, r) in enumerate(zip(acqs, res)):
            b = ol.RefAcqBlock(cshort=int(item == "16ic"), prn=(PRNS + [31])[k], **cfg)
            b.set_local_code(codes[k])
            b.start(777 + 500)
            seg = host[500:4500] if item == "32fc" else host[1000:9000]
            st, ev = b.dwell(seg)
            assert {1: 1, 0: 2}[r.decision] == ev, (item, k, r.decision, ev)
            assert abs(r.input_power - st.input_power) <= 1e-4 * st.input_power
            if k < 8:
                assert (r.acq_delay_samples, r.acq_doppler_hz, r.acq_samplestamp) == (st.acq_delay_samples, st.acq_doppler_hz, st.acq_samplestamp_samples), (item, k)
                assert abs(r.test_statistic - st.test_statistics) <= 1e-4 * st.test_statistics and ev == 1
            b.close()
            a.close()


@pytest.mark.parametrize("order", [2, 3])
def test_kf_loop_on_b200_matches_oracle_loop_and_the_reference_block(service, order):
    """gps_l1_ca_kf_tracking_cc: the Kalman carrier loop mirror (host/gb200_kf_tracking.hpp) closed on the B200
    real-code correlator -- configured like the reference object the block owns: high-dynamics code resampler, plain rotator --
    vs the same loop on the oracle correlator AND vs the reference block itself (oracle/_ref).  Block boundaries and flags equal;
    Doppler within 0.01 Hz, prompts within 1e-4 of the peak in the median, code phase within 1e-3 samples over 600 epochs."""
    import gnss_b200 as gb
    import oracle_lib as ol
    import test_host_cpu as thc
    from test_ref_blocks_cpu import quiet

    fs, doppler, delay = 4e6, -2650.0, 912.3
    x, code = thc.caid_scene(640, doppler, delay, cn0=47.0, prn=23, seed=20260033)
    service.push(2, x, gb.ITEM_32FC, 0)
    kw = dict(fs_in=fs, vector_length=4000, dll_bw_hz=2.0, early_late_space_chips=0.5, order=order)
    args = dict(code_f32=code.real, acq_delay=delay + 0.4, acq_doppler=doppler + 45.0, acq_stamp=0, acq_doppler_step=250, n_steps=600)
    ref = hl.run_oracle_kf(x, 0, **args, **kw)
    got = hl.run_b200_kf(service, 2, gb.ITEM_32FC, 0, **args, **kw)
    with quiet():
        b = ol.RefSiblingTracker(ol.REF_TRK_KF, fs, 4000, 0.0, 2.0, early_late_space_chips=0.5, order=order)
        b.start(23, delay + 0.4, doppler + 45.0, 0, 0, acq_doppler_step=250)
        blk = b.run(x, 600)
        b.close()
    F = {n: i for i, n in enumerate(hl.KF_EPOCH_FIELDS)}
    assert len(blk) == 600 and np.array_equal(ref[1:, :10], blk[1:, :10])  # mirror on the oracle correlator == block (also a CPU test)
    assert np.array_equal(got[:, F["prn_length_samples"]], blk[:, 11])
    assert np.array_equal(got[:, F["tracking_sample_counter"]], ref[:, F["tracking_sample_counter"]])
    assert np.array_equal(got[:, F["flags"]], ref[:, F["flags"]]) and np.all(got[-300:, F["flags"]] == 1)
    rel = np.abs(got[1:, F["abs_p"]] - blk[1:, 8]) / blk[1:, 8].max()
    dd = np.abs(got[:, F["carrier_doppler_hz"]] - blk[:, 3])
    print("KF order %d closed loop vs the block: doppler diff max %.3g Hz, prompt magnitude diff max %.3g of peak, final doppler %.1f Hz" % (
        order, dd[1:].max(), rel.max(), got[-1, F["carrier_doppler_hz"]]))
    # the two closed loops drift apart by ~1e-6 samples of code phase, so now and then ONE sample of one epoch falls on the other side
    # of a chip boundary and moves that epoch's prompt by up to 2 |x| (measured: one epoch of 600 at 3.2e-3 of the peak, all others
    # below 5e-5; the same correlate call with the block's own arguments agrees to 2e-5): median tight, maximum = one flipped chip
    assert dd[1:].max() < 0.01 and np.median(rel) < 1e-4 and rel.max() < 1.5e-2
    assert abs(got[-200:, F["carrier_doppler_hz"]].mean() - doppler) < 3.0
    assert np.abs(got[1:, F["code_phase_samples"]] - blk[1:, 4]).max() < 1e-3
