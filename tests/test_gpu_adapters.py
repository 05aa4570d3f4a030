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
