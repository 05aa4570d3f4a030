"""The outer drop-in boundary: the *_B200 adapter classes of gnss-sdr-1_b200/adapters -- AcquisitionInterface /
TrackingInterface implementations made by name through the factory entries, configured by the reference's own .conf keys -- driven
the way Channel, ChannelFsm and the GNU Radio scheduler drive them (tests/native/adapter_harness.cc), on the B200.

The configuration is the reference's conf/gnss-sdr_GPS_L1_gr_complex.conf with only the two implementation= lines changed; it is
parsed by the reference's FileConfiguration / INIReader (compiled from /root/reference into the test library)."""
import ctypes as C
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "native", "libb200_adapters_test.so")


def harness():
    if not os.path.exists(LIB):
        pytest.skip("tests/native/libb200_adapters_test.so not built (needs /root/reference at build time)")
    L = C.CDLL(LIB)
    L.b200_adapter_receiver.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong, C.c_int, C.POINTER(C.c_uint32), C.c_char_p, C.c_int, C.POINTER(C.c_double), C.c_char_p, C.c_int]
    return L


def run_receiver(conf, samples, n_items, prns, system_signal, chunk=8192):
    L = harness()
    out = (C.c_double * (16 * len(prns)))()
    names = C.create_string_buffer(256)
    rc = L.b200_adapter_receiver(conf.encode(), samples.ctypes.data_as(C.c_void_p), n_items, len(prns), (C.c_uint32 * len(prns))(*prns), system_signal, chunk, out, names, 256)
    return rc, np.array(out[:]).reshape(len(prns), 16), names.value.decode()


def test_reference_conf_with_b200_implementations_acquires_and_tracks_c1(golden_dir):
    """8 channels (BASELINE C1) on a 4 Msps gr_complex scene: every channel's acquisition adapter finds its satellite at the scene's
    delay / Doppler bin, hands over to its tracking adapter through the Gnss_Synchro, and the tracking adapter locks: Doppler on the
    truth, CN0 near the simulated one, no loss of lock, one Gnss_Synchro per millisecond."""
    prns = [1, 3, 6, 9, 12, 17, 22, 28]
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=50 - 0.5 * k, doppler_hz=-4230.0 + 1250.0 * k, delay_samples=137 + 487 * k) for k, p in enumerate(prns)]  # 20 Hz off the conf's 250 Hz bins
    fs, n = 4e6, 4_800_000  # 1.2 s
    x = synth.make_iq(fs, n, sats, 20260001)
    rc, out, names = run_receiver(golden_dir + "/gnss-sdr_GPS_L1_gr_complex_B200.conf", x, n, prns, b"G1C")
    assert rc == 0, (rc, names)
    assert names == "GPS_L1_CA_PCPS_Acquisition_B200 GPS_L1_CA_DLL_PLL_Tracking_B200"
    for k, s in enumerate(sats):
        o = out[k]
        assert o[0] >= 1, (k, o)                                        # a positive acquisition
        assert o[4] > 0 and o[4] % 1 == 0                               # Acq_samplestamp_samples = end of the searched block
        want = (s["delay_samples"] - (o[4] - 4000)) % 4000              # code start relative to the first sample of that block
        assert min(abs(o[2] - want), 4000 - abs(o[2] - want)) <= 1.0, (k, o[2], want)  # Acq_delay_samples
        assert abs(o[3] - s["doppler_hz"]) <= 250.0, (k, o[3])          # nearest 250 Hz bin of the conf's +-10 kHz grid
        assert o[8] == 0, (k, "loss of lock", o)
        # one Gnss_Synchro per millisecond from the hand-over to the end of the stream (the weak satellites need a few attempts
        # at the conf's threshold 0.008: negative acquisitions restart the search, as ChannelFsm does)
        # (the standby tracking block runs ahead of a retrying acquisition block in this scheduler, so the count is not n - stamp)
        assert o[5] > 300 and o[9] > n - 3 * 4000, (k, o[5], o[4], o[1], o[9])
        assert abs(o[6] - s["doppler_hz"]) < 5.0, (k, o[6], s["doppler_hz"])
        assert abs(o[7] - s["cn0_dbhz"]) < 5.0, (k, o[7])  # a 20-prompt SNV estimate under 7 interfering satellites
        assert o[11] == 8 and o[12] == 8                                # item_size() of gr_complex


def test_galileo_e1_adapters_with_cshort_items(tmp_path):
    """Galileo_E1_PCPS_Ambiguous_Acquisition_B200 + Galileo_E1_DLL_PLL_VEML_Tracking_B200 from a written .conf: item_type=cshort (the
    stream stays packed int16 on the device), 8 Msps, pilot tracking with the data-prompt tap; per-role keys and E1 defaults."""
    conf = tmp_path / "e1.conf"
    conf.write_text("""[GNSS-SDR]
GNSS-SDR.internal_fs_sps=8000000
GNSS-SDR.b200_batch_window_us=0
Acquisition_1B.implementation=Galileo_E1_PCPS_Ambiguous_Acquisition_B200
Acquisition_1B.item_type=cshort
Acquisition_1B.threshold=0.0012
Acquisition_1B.doppler_max=5000
Acquisition_1B.doppler_step=125
Tracking_1B.implementation=Galileo_E1_DLL_PLL_VEML_Tracking_B200
Tracking_1B.item_type=cshort
Tracking_1B.track_pilot=true
Tracking_1B.pll_bw_hz=15.0
Tracking_1B.dll_bw_hz=2.0
Tracking_1B.pll_filter_order=2
""")
    prns = [11, 19]
    sats = [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=p, cn0_dbhz=50 - 2 * k, doppler_hz=1270.0 - 2875.0 * k, delay_samples=9123 + 7001 * k,
                 secondary="0011100000001010110110010") for k, p in enumerate(prns)]
    fs, n = 8e6, 5_600_000  # 0.7 s
    x = synth.make_iq(fs, n, sats, 20260002)
    q = synth.quantize(x, "16ic")
    rc, out, names = run_receiver(str(conf), q, n, prns, b"E1B", chunk=65536)
    assert rc == 0, (rc, names)
    for k, s in enumerate(sats):
        o = out[k]
        want = (s["delay_samples"] - (o[4] - 32000)) % 32000
        assert o[0] >= 1 and min(abs(o[2] - want), 32000 - abs(o[2] - want)) <= 1.0 and abs(o[3] - s["doppler_hz"]) <= 125.0, (k, o, want)
        assert o[8] == 0 and o[5] > 140, (k, o)                         # 4 ms epochs
        assert abs(o[6] - s["doppler_hz"]) < 3.0, (k, o[6])
        assert o[11] == 4 and o[12] == 4                                # item_size() of cshort
    # an unknown implementation name is not ours: the factory returns nullptr and the reference's chain would go on
    bad = tmp_path / "bad.conf"
    bad.write_text(conf.read_text().replace("Galileo_E1_PCPS_Ambiguous_Acquisition_B200", "Galileo_E1_PCPS_Ambiguous_Acquisition"))
    rc, _, _ = run_receiver(str(bad), q, n, prns, b"E1B")
    assert rc == -1


# ------------------------------------------------------------------------------------------------ table-driven adapters (b200_signal_adapters.*)
def _conf(tmp_path, name, text):
    p = tmp_path / name
    p.write_text("[GNSS-SDR]\nGNSS-SDR.b200_batch_window_us=0\n" + text)
    return str(p)


def _check(out, sats, prns, n_code, dstep, min_epochs, dopp_tol=4.0, cn0_tol=None):
    """cn0_tol None: only "above the lock detector's floor".  On signals with a secondary code the reference block's own CN0 estimate
    sits 10-17 dB under the scene's (tests/test_ref_blocks_cpu.py::test_other_signals_mirror_equals_the_tracking_block holds the
    mirror to the block there), so it is not compared with the truth."""
    for k, s in enumerate(sats):
        o = out[k]
        want = (s["delay_samples"] - (o[4] - n_code)) % n_code
        assert o[0] >= 1 and min(abs(o[2] - want), n_code - abs(o[2] - want)) <= 1.0, (k, o, want)
        assert abs(o[3] - s["doppler_hz"]) <= dstep, (k, o[3])
        if min_epochs:
            assert o[8] == 0 and o[5] > min_epochs, (k, o)
            assert abs(o[6] - s["doppler_hz"]) < dopp_tol, (k, o[6], s["doppler_hz"])
            assert (abs(o[7] - s["cn0_dbhz"]) < cn0_tol) if cn0_tol else (o[7] > 25.0), (k, o[7])


def test_gps_l5_adapters_from_the_signal_table(tmp_path):
    """GPS_L5i_PCPS_Acquisition_B200 + GPS_L5_DLL_PLL_Tracking_B200 (BASELINE C3's signal) made by name through the factory: the L5i
    adapter's keys and derivations (floor code length, samples_per_code from the period, per-channel pfa lookup) and the L5 tracker's
    defaults; L5I data component with its 10-bit Neuman-Hofman code, 25 Msps gr_complex (C3's rate: 2.4 samples per chip, so the
    one-sample ambiguity of the acquisition is inside the DLL's pull-in range), 0.3 s."""
    conf = _conf(tmp_path, "l5.conf", """GNSS-SDR.internal_fs_sps=25000000
Acquisition_L5.implementation=GPS_L5i_PCPS_Acquisition_B200
Acquisition_L5.threshold=0.002
Acquisition_L5.doppler_max=5000
Acquisition_L5.doppler_step=250
Tracking_L5.implementation=GPS_L5_DLL_PLL_Tracking_B200
Tracking_L5.pll_bw_hz=30.0
Tracking_L5.dll_bw_hz=2.0
""")
    prns = [3, 24]
    sats = [dict(kind="gps_l5i", prn=p, cn0_dbhz=50 - 2 * k, doppler_hz=1520.0 - 2730.0 * k, delay_samples=2345 + 5001 * k, secondary="0000110101") for k, p in enumerate(prns)]
    fs, n = 25e6, 7_500_000
    x = synth.make_iq(fs, n, sats, 20260061)
    rc, out, names = run_receiver(conf, x, n, prns, b"GL5", chunk=50000)
    assert rc == 0 and names == "GPS_L5i_PCPS_Acquisition_B200 GPS_L5_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 25000, 250.0, 180)


def test_galileo_e5a_adapters_from_the_signal_table(tmp_path):
    """Galileo_E5a_Pcps_Acquisition_B200 + Galileo_E5a_DLL_PLL_Tracking_B200: the E5a adapter's own rules (sampled_ms fixed at 1,
    first/second-peak statistic by default, code length recomputed with round(), "5I" replica) and the E5a tracker defaults
    (PLL 20 Hz / DLL 20 Hz); E5a-I primary code from the reference generator (oracle/_ref) with the 20-bit secondary code."""
    import oracle_lib as ol

    R = ol.ref()
    if R is None:
        pytest.skip("oracle/_ref not built")
    conf = _conf(tmp_path, "e5a.conf", """GNSS-SDR.internal_fs_sps=25000000
Acquisition_5X.implementation=Galileo_E5a_Pcps_Acquisition_B200
Acquisition_5X.threshold=2.5
Acquisition_5X.doppler_max=5000
Acquisition_5X.doppler_step=250
Tracking_5X.implementation=Galileo_E5a_DLL_PLL_Tracking_B200
Tracking_5X.dll_bw_hz=5.0
""")
    prns = [11, 30]
    sats = []
    for k, p in enumerate(prns):
        c = np.zeros(2 * 10230, np.float32)
        R.gbref_galileo_e5a_code_primary(c, p, b"5I")
        sats.append(dict(kind="e5a_i", prn=p, table=c[0::2], table_rate=10.23e6, carrier_hz=1176.45e6, cn0_dbhz=51 - 2 * k, doppler_hz=-2243.0 + 3490.0 * k,  # within 7 Hz of a bin: a Costas loop on the 20-bit secondary code false-locks 25 Hz off otherwise
                        
                         delay_samples=777 + 6007 * k, secondary="10000100001011101001"))
    fs, n = 25e6, 7_500_000
    x = synth.make_iq(fs, n, sats, 20260062)
    rc, out, names = run_receiver(conf, x, n, prns, b"E5X", chunk=50000)
    assert rc == 0 and names == "Galileo_E5a_Pcps_Acquisition_B200 Galileo_E5a_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 25000, 250.0, 180)


def test_beidou_b1i_adapters_from_the_signal_table(tmp_path):
    """BEIDOU_B1I_PCPS_Acquisition_B200 + BEIDOU_B1I_DLL_PLL_Tracking_B200 (BASELINE C4's third constellation): the older adapter
    layout (round() code length, samples_per_ms = samples_per_code = code length, Acq_Conf::ms_per_code left at its default), cbyte
    items kept packed on the device; MEO satellites with the 20-bit Neuman-Hofman code."""
    conf = _conf(tmp_path, "b1i.conf", """GNSS-SDR.internal_fs_sps=8000000
Acquisition_B1.implementation=BEIDOU_B1I_PCPS_Acquisition_B200
Acquisition_B1.item_type=cbyte
Acquisition_B1.threshold=0.003
Acquisition_B1.doppler_max=5000
Acquisition_B1.doppler_step=250
Tracking_B1.implementation=BEIDOU_B1I_DLL_PLL_Tracking_B200
Tracking_B1.item_type=cbyte
Tracking_B1.pll_bw_hz=30.0
""")
    prns = [8, 14]
    sats = [dict(kind="bds_b1i", prn=p, cn0_dbhz=51 - 2 * k, doppler_hz=2480.0 - 3210.0 * k, delay_samples=1234 + 377 * k, secondary="00000100110101001110") for k, p in enumerate(prns)]
    # (small code delays on purpose: with Acq_Conf::ms_per_code left at 0 the block zero-pads to twice the length and correlates linearly,
    # so a satellite whose code starts late in the block overlaps the replica only partly -- the reference adapter's own sensitivity loss)
    fs, n = 8e6, 4_000_000
    x = synth.make_iq(fs, n, sats, 20260063)
    q = synth.quantize(x, "8ic")
    rc, out, names = run_receiver(conf, q, n, prns, b"CB1", chunk=16000)
    assert rc == 0 and names == "BEIDOU_B1I_PCPS_Acquisition_B200 BEIDOU_B1I_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 8000, 250.0, 300)
    assert out[0][11] == 2 and out[0][12] == 2  # item_size() of cbyte


def test_glonass_l1_acquisition_adapter_adds_the_fdma_offset(tmp_path):
    """GLONASS_L1_CA_PCPS_Acquisition_B200 (no Tracking_1G line: acquisition only): every satellite is searched on its own FDMA
    channel -- the block reads Gnss_Synchro::Signal == "1G" and the slot's channel number (GLONASS_PRN) in set_local_code and adds
    k * 562.5 kHz to every Doppler bin, as pcps_acquisition::is_fdma does."""
    conf = _conf(tmp_path, "glo.conf", """GNSS-SDR.internal_fs_sps=8000000
Acquisition_1G.implementation=GLONASS_L1_CA_PCPS_Acquisition_B200
Acquisition_1G.threshold=0.003
Acquisition_1G.doppler_max=5000
Acquisition_1G.doppler_step=250
""")
    prns = [1, 2, 3]          # GLONASS_PRN: slots 1, 2, 3 -> channels 1, -4, 5
    chans = [1, -4, 5]
    sats = [dict(kind="glo_l1_ca", prn=1, cn0_dbhz=50 - k, doppler_hz=-1730.0 + 1500.0 * k, carrier_offset_hz=562500.0 * ch, delay_samples=431 + 2003 * k)
            for k, ch in enumerate(chans)]
    fs, n = 8e6, 400_000
    x = synth.make_iq(fs, n, sats, 20260064)
    rc, out, names = run_receiver(conf, x, n, prns, b"R1G", chunk=16000)
    assert rc == 0 and names.strip() == "GLONASS_L1_CA_PCPS_Acquisition_B200", (rc, names)
    _check(out, sats, prns, 8000, 250.0, 0)


# ------------------------------------------------------------------------------------------------ the older per-signal trackers (b200_sibling_tracking*)
@pytest.mark.parametrize("impl,item", [("GPS_L1_CA_KF_Tracking_B200", "gr_complex"), ("GPS_L1_CA_DLL_PLL_C_Aid_Tracking_B200", "gr_complex"),
                                       ("GPS_L1_CA_DLL_PLL_C_Aid_Tracking_B200", "cshort")])
def test_gps_l1_sibling_tracker_adapters(tmp_path, impl, item):
    """GPS_L1_CA_KF_Tracking_B200 and GPS_L1_CA_DLL_PLL_C_Aid_Tracking_B200 (gr_complex -> the _cc loop, cshort -> the _sc loop on the
    int16 correlator, as the reference adapter chooses) behind the GPS L1 acquisition adapter: made by name, configured by the
    reference adapters' keys, one Gnss_Synchro per general_work, locked on the scene's Doppler."""
    conf = _conf(tmp_path, "sib.conf", """GNSS-SDR.internal_fs_sps=4000000
Acquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_B200
Acquisition_1C.item_type=%s
Acquisition_1C.threshold=0.008
Acquisition_1C.doppler_max=5000
Acquisition_1C.doppler_step=250
Tracking_1C.implementation=%s
Tracking_1C.item_type=%s
Tracking_1C.pll_bw_hz=35.0
Tracking_1C.dll_bw_hz=2.0
Tracking_1C.order=2
""" % (item, impl, item))
    prns = [5, 20]
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=49 - k, doppler_hz=-1490.0 + 2755.0 * k, delay_samples=911 + 1501 * k) for k, p in enumerate(prns)]
    fs, n = 4e6, 3_200_000  # 0.8 s
    x = synth.make_iq(fs, n, sats, 20260081)
    samples = x if item == "gr_complex" else np.ascontiguousarray(np.rint(np.stack([x.real, x.imag], -1) * 4.0).astype(np.int16).reshape(-1))
    rc, out, names = run_receiver(conf, samples, n, prns, b"G1C")
    assert rc == 0 and names == "GPS_L1_CA_PCPS_Acquisition_B200 " + impl, (rc, names)
    # the Kalman tracker has no frequency pull-in aid: its Doppler state closes in over seconds, not within this 0.8 s stream
    _check(out, sats, prns, 4000, 250.0, 500, dopp_tol=12.0 if "KF" in impl else 5.0, cn0_tol=6.0)
    assert out[0][12] == (8 if item == "gr_complex" else 4)


def test_glonass_l1_acquisition_and_tracking_adapters(tmp_path):
    """GLONASS_L1_CA_PCPS_Acquisition_B200 + GLONASS_L1_CA_DLL_PLL_Tracking_B200 on three FDMA channels: the tracker takes the slot's
    channel number from GLONASS_PRN and tracks carrier = Doppler + k * 562.5 kHz with the complex-code correlator."""
    conf = _conf(tmp_path, "glo2.conf", """GNSS-SDR.internal_fs_sps=8000000
Acquisition_1G.implementation=GLONASS_L1_CA_PCPS_Acquisition_B200
Acquisition_1G.threshold=0.003
Acquisition_1G.doppler_max=5000
Acquisition_1G.doppler_step=250
Tracking_1G.implementation=GLONASS_L1_CA_DLL_PLL_Tracking_B200
Tracking_1G.pll_bw_hz=40.0
Tracking_1G.dll_bw_hz=2.0
""")
    # slots 1, 11, 12 -> channels 1, 0, -1.  (On the far channels the reference's float32-accumulated wipe-off phase -- 2 rad per sample at
    # k = -4 -- drifts enough to move the peak one 250 Hz bin; the engine reproduces that, tests/test_gpu_configs.py, and a tracker
    # started 230 Hz off does not pull in, there or here.)
    prns, chans = [1, 11, 12], [1, 0, -1]
    sats = [dict(kind="glo_l1_ca", prn=1, cn0_dbhz=50 - k, doppler_hz=-1730.0 + 1500.0 * k, carrier_offset_hz=562500.0 * ch, delay_samples=431 + 2003 * k)
            for k, ch in enumerate(chans)]
    fs, n = 8e6, 4_000_000  # 0.5 s
    x = synth.make_iq(fs, n, sats, 20260082)
    rc, out, names = run_receiver(conf, x, n, prns, b"R1G", chunk=16000)
    assert rc == 0 and names == "GLONASS_L1_CA_PCPS_Acquisition_B200 GLONASS_L1_CA_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 8000, 250.0, 300, dopp_tol=5.0, cn0_tol=6.0)


@pytest.mark.parametrize("impl,keys", [
    ("GPS_L1_CA_PCPS_Tong_Acquisition_B200", "Acquisition_1C.threshold=0.008\nAcquisition_1C.tong_init_val=1\nAcquisition_1C.tong_max_val=2\n"),
    ("GPS_L1_CA_PCPS_QuickSync_Acquisition_B200", "Acquisition_1C.threshold=0.1\nAcquisition_1C.folding_factor=2\nAcquisition_1C.coherent_integration_time_ms=4\n"),
    ("GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200", "Acquisition_1C.threshold=2.5\nAcquisition_1C.max_dwells=2\n"),
])
def test_gps_l1_sibling_acquisition_adapters(tmp_path, impl, keys):
    """The sibling GPS L1 C/A acquisition blocks as adapters made by name (Tong detector, QuickSync, fine Doppler), each handing over to
    the GPS L1 DLL/PLL tracking adapter through the channel's Gnss_Synchro: right delay, Doppler within a bin (the fine-Doppler one
    within 20 Hz), tracking locked for the rest of the stream."""
    conf = _conf(tmp_path, "sibacq.conf", """GNSS-SDR.internal_fs_sps=4000000
Acquisition_1C.implementation=%s
Acquisition_1C.doppler_max=5000
Acquisition_1C.doppler_step=250
%sTracking_1C.implementation=GPS_L1_CA_DLL_PLL_Tracking_B200
Tracking_1C.pll_bw_hz=35.0
Tracking_1C.dll_bw_hz=2.0
""" % (impl, keys))
    prns = [5, 20]
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=50 - k, doppler_hz=-1490.0 + 2755.0 * k, delay_samples=911 + 1501 * k) for k, p in enumerate(prns)]
    fs, n = 4e6, 3_200_000  # 0.8 s
    x = synth.make_iq(fs, n, sats, 20260083)
    rc, out, names = run_receiver(conf, x, n, prns, b"G1C", chunk=65536)
    assert rc == 0 and names == impl + " GPS_L1_CA_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 4000, 20.0 if "Fine" in impl else 250.0, 500, dopp_tol=5.0, cn0_tol=6.0)


def test_galileo_e1_8ms_acquisition_adapter(tmp_path):
    """Galileo_E1_PCPS_8ms_Ambiguous_Acquisition_B200 (8 ms dwell, codes A and B searched in one batch, coherent_integration_time_ms a
    multiple of 4) handing over to the Galileo E1 VEML tracking adapter."""
    conf = _conf(tmp_path, "e1_8ms.conf", """GNSS-SDR.internal_fs_sps=4000000
Acquisition_1B.implementation=Galileo_E1_PCPS_8ms_Ambiguous_Acquisition_B200
Acquisition_1B.coherent_integration_time_ms=8
Acquisition_1B.threshold=0.002
Acquisition_1B.doppler_max=5000
Acquisition_1B.doppler_step=125
Tracking_1B.implementation=Galileo_E1_DLL_PLL_VEML_Tracking_B200
Tracking_1B.track_pilot=true
Tracking_1B.pll_bw_hz=15.0
Tracking_1B.dll_bw_hz=2.0
Tracking_1B.pll_filter_order=2
""")
    prns = [11, 19]
    sats = [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=p, cn0_dbhz=50 - 2 * k, doppler_hz=1270.0 - 2875.0 * k, delay_samples=9123 + 3001 * k,
                 secondary="0011100000001010110110010") for k, p in enumerate(prns)]
    fs, n = 4e6, 2_800_000  # 0.7 s
    x = synth.make_iq(fs, n, sats, 20260084)
    rc, out, names = run_receiver(conf, x, n, prns, b"E1B", chunk=65536)
    assert rc == 0 and names == "Galileo_E1_PCPS_8ms_Ambiguous_Acquisition_B200 Galileo_E1_DLL_PLL_VEML_Tracking_B200", (rc, names)
    for k, s in enumerate(sats):
        o = out[k]
        want = (s["delay_samples"] - (o[4] - 32000)) % 16000
        assert o[0] >= 1 and min(abs(o[2] - want), 16000 - abs(o[2] - want)) <= 1.0 and abs(o[3] - s["doppler_hz"]) <= 125.0, (k, o, want)
        assert o[8] == 0 and o[5] > 120 and abs(o[6] - s["doppler_hz"]) < 3.0, (k, o)


def test_galileo_e5a_noncoherent_iq_caf_acquisition_adapter(tmp_path):
    """Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200 (I and Q replicas, non-coherent combination on the device, CAF window) in front of
    the Galileo E5a tracking adapter; E5a-I + E5a-Q in quadrature from the reference generator (oracle/_ref)."""
    import oracle_lib as ol

    R = ol.ref()
    if R is None:
        pytest.skip("oracle/_ref not built")
    conf = _conf(tmp_path, "e5a_caf.conf", """GNSS-SDR.internal_fs_sps=25000000
Channel.signal=5X
Acquisition_5X.implementation=Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200
Acquisition_5X.threshold=0.001
Acquisition_5X.doppler_max=5000
Acquisition_5X.doppler_step=250
Acquisition_5X.CAF_window_hz=0
Acquisition_5X.max_dwells=1
Tracking_5X.implementation=Galileo_E5a_DLL_PLL_Tracking_B200
Tracking_5X.dll_bw_hz=5.0
""")
    prns = [11, 30]
    fs, n = 25e6, 7_500_000
    x = None
    sats = []
    for k, p in enumerate(prns):
        ci = np.zeros(2 * 10230, np.float32)
        cq = np.zeros(2 * 10230, np.float32)
        R.gbref_galileo_e5a_code_primary(ci, p, b"5I")
        R.gbref_galileo_e5a_code_primary(cq, p, b"5Q")
        base = dict(prn=p, table_rate=10.23e6, carrier_hz=1176.45e6, cn0_dbhz=48 - 2 * k, doppler_hz=-2243.0 + 3490.0 * k, delay_samples=777 + 6007 * k)
        s_i = dict(kind="e5a_i", table=ci[0::2], secondary="10000100001011101001", phase=0.0, **base)
        s_q = dict(kind="e5a_q", table=cq[1::2], phase=np.pi / 2, **base)  # the pilot rides on the quadrature carrier
        part = synth.make_iq(fs, n, [s_i, s_q], 20260085 + k, noise=(k == 0))
        x = part if x is None else x + part
        sats.append(dict(cn0_dbhz=51 - 2 * k, **{kk: base[kk] for kk in ("doppler_hz", "delay_samples")}))
    rc, out, names = run_receiver(conf, np.ascontiguousarray(x, np.complex64), n, prns, b"E5X", chunk=50000)
    assert rc == 0 and names == "Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200 Galileo_E5a_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 25000, 250.0, 180)


def _ref_table(fn_name, prn, length, *extra):
    import oracle_lib as ol

    R = ol.ref()
    if R is None:
        pytest.skip("oracle/_ref not built")
    fn = getattr(R, fn_name)
    fn.argtypes = [np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")] + [C.c_int if fn_name.startswith("gbref_beidou") else C.c_uint] + [C.c_uint] * len(extra)
    fn.restype = None
    c = np.zeros(length, np.float32)
    fn(c, prn, *extra)
    return c


def test_gps_l2c_adapters_from_the_signal_table(tmp_path):
    """GPS_L2_M_PCPS_Acquisition_B200 + GPS_L2_M_DLL_PLL_Tracking_B200: 20 ms code (ms_per_code = 20 is also the default dwell), 80000-point
    search at 4 Msps, 20 ms tracking epochs without secondary code; L2 CM code from the reference generator (oracle/_ref)."""
    conf = _conf(tmp_path, "l2c.conf", """GNSS-SDR.internal_fs_sps=4000000
Acquisition_2S.implementation=GPS_L2_M_PCPS_Acquisition_B200
Acquisition_2S.threshold=0.0008
Acquisition_2S.doppler_max=2000
Acquisition_2S.doppler_step=25
Tracking_2S.implementation=GPS_L2_M_DLL_PLL_Tracking_B200
Tracking_2S.pll_bw_hz=4.0
Tracking_2S.dll_bw_hz=0.75
""")
    prns = [7, 15]
    sats = [dict(kind="l2cm", prn=p, table=_ref_table("gbref_gps_l2c_m_code_float", p, 10230), table_rate=0.5115e6, carrier_hz=1227.6e6, cn0_dbhz=44 - 2 * k,
                 doppler_hz=-1004.0 + 1530.0 * k, delay_samples=31234 + 20011 * k) for k, p in enumerate(prns)]  # within 4 Hz of a 25 Hz bin: a 20 ms PLL pulls in +-12 Hz
    fs, n = 4e6, 6_000_000  # 1.5 s
    x = synth.make_iq(fs, n, sats, 20260091)
    rc, out, names = run_receiver(conf, x, n, prns, b"G2S", chunk=160000)
    assert rc == 0 and names == "GPS_L2_M_PCPS_Acquisition_B200 GPS_L2_M_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 80000, 25.0, 40, dopp_tol=3.0)


def test_beidou_b3i_adapters_from_the_signal_table(tmp_path):
    """BEIDOU_B3I_PCPS_Acquisition_B200 + BEIDOU_B3I_DLL_PLL_Tracking_B200 at 25 Msps (10.23 Mcps code, 20-bit Neuman-Hofman code)."""
    conf = _conf(tmp_path, "b3i.conf", """GNSS-SDR.internal_fs_sps=25000000
Acquisition_B3.implementation=BEIDOU_B3I_PCPS_Acquisition_B200
Acquisition_B3.threshold=0.0012
Acquisition_B3.doppler_max=5000
Acquisition_B3.doppler_step=250
Tracking_B3.implementation=BEIDOU_B3I_DLL_PLL_Tracking_B200
Tracking_B3.pll_bw_hz=30.0
""")
    prns = [8, 14]
    sats = [dict(kind="b3i", prn=p, table=_ref_table("gbref_beidou_b3i_code_float", p, 10230, 0), table_rate=10.23e6, carrier_hz=1268.52e6, cn0_dbhz=54 - 2 * k,
                 doppler_hz=2480.0 - 3210.0 * k, delay_samples=1234 + 977 * k, secondary="00000100110101001110") for k, p in enumerate(prns)]
    fs, n = 25e6, 7_500_000
    x = synth.make_iq(fs, n, sats, 20260092)
    rc, out, names = run_receiver(conf, x, n, prns, b"CB3", chunk=50000)
    assert rc == 0 and names == "BEIDOU_B3I_PCPS_Acquisition_B200 BEIDOU_B3I_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 25000, 250.0, 180)


def test_glonass_l2_acquisition_and_tracking_adapters(tmp_path):
    """GLONASS_L2_CA_PCPS_Acquisition_B200 + GLONASS_L2_CA_DLL_PLL_Tracking_B200: signal "2G", channel spacing 437.5 kHz (DFRQ2_GLO)."""
    conf = _conf(tmp_path, "glo_l2.conf", """GNSS-SDR.internal_fs_sps=8000000
Acquisition_2G.implementation=GLONASS_L2_CA_PCPS_Acquisition_B200
Acquisition_2G.threshold=0.003
Acquisition_2G.doppler_max=5000
Acquisition_2G.doppler_step=250
Tracking_2G.implementation=GLONASS_L2_CA_DLL_PLL_Tracking_B200
Tracking_2G.pll_bw_hz=40.0
Tracking_2G.dll_bw_hz=2.0
""")
    prns, chans = [1, 11, 12], [1, 0, -1]
    sats = [dict(kind="glo_l1_ca", prn=1, cn0_dbhz=50 - k, doppler_hz=-1730.0 + 1500.0 * k, carrier_offset_hz=437500.0 * ch, delay_samples=431 + 2003 * k)
            for k, ch in enumerate(chans)]  # the L2 C/A code is the L1 C/A code
    fs, n = 8e6, 4_000_000
    x = synth.make_iq(fs, n, sats, 20260093)
    rc, out, names = run_receiver(conf, x, n, prns, b"R2G", chunk=16000)
    assert rc == 0 and names == "GLONASS_L2_CA_PCPS_Acquisition_B200 GLONASS_L2_CA_DLL_PLL_Tracking_B200", (rc, names)
    _check(out, sats, prns, 8000, 250.0, 300, dopp_tol=5.0, cn0_tol=6.0)
