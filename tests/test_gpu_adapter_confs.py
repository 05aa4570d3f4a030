"""'Same keys, same defaults' checked against the reference's OWN adapters (outer boundary).

tests/native/adapter_conf_compare.cc compiles src/algorithms/{acquisition,tracking}/adapters/*.cc unmodified (stand-in GNU Radio headers;
pcps_acquisition / dll_pll_veml_tracking from oracle/_ref), constructs the reference adapter and the *_B200 adapter on the same
FileConfiguration, and hands back the Acq_Conf / Dll_Pll_Conf found in the BLOCK each adapter made (plus the threshold it set and
item_size()).  Every value must be equal -- for an empty role (all defaults), for a role with every key set, and for the odd cases
(bit transition, resampler, pfa, extend_correlation_symbols clamps, pilot requests on signals without one).  A GPU is needed only
because the *_B200 adapters open the device service in their constructors."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "native", "libb200_adapter_conf_test.so")

ACQ_FIELDS = ("sampled_ms", "ms_per_code", "samples_per_chip", "max_dwells", "doppler_max", "num_doppler_bins_step2", "doppler_step2", "fs_in", "samples_per_ms",
              "samples_per_code", "bit_transition_flag", "use_CFAR_algorithm_flag", "dump", "blocking", "blocking_on_standby", "make_2_steps", "dump_channel",
              "use_automatic_resampler", "resampler_ratio", "resampled_fs", "threshold", "item_size")
TRK_FIELDS = ("fs_in", "vector_length", "dump", "dump_mat", "pll_bw_hz", "dll_bw_hz", "fll_bw_hz", "pll_bw_narrow_hz", "dll_bw_narrow_hz", "early_late_space_chips",
              "very_early_late_space_chips", "early_late_space_narrow_chips", "very_early_late_space_narrow_chips", "extend_correlation_symbols", "high_dyn",
              "cn0_samples", "cn0_min", "max_lock_fail", "carrier_lock_th", "track_pilot", "system", "signal0", "signal1", "pll_filter_order", "dll_filter_order",
              "fll_filter_order", "enable_fll_pull_in", "pull_in_time_s", "smoother_length", "item_size")

ACQ = [  # (role, reference implementation)
    ("Acquisition_1C", "GPS_L1_CA_PCPS_Acquisition"), ("Acquisition_1B", "Galileo_E1_PCPS_Ambiguous_Acquisition"), ("Acquisition_L5", "GPS_L5i_PCPS_Acquisition"),
    ("Acquisition_5X", "Galileo_E5a_Pcps_Acquisition"), ("Acquisition_B1", "BEIDOU_B1I_PCPS_Acquisition"), ("Acquisition_1G", "GLONASS_L1_CA_PCPS_Acquisition"),
    ("Acquisition_2S", "GPS_L2_M_PCPS_Acquisition"), ("Acquisition_B3", "BEIDOU_B3I_PCPS_Acquisition"), ("Acquisition_2G", "GLONASS_L2_CA_PCPS_Acquisition"),
]
TRK = [
    ("Tracking_1C", "GPS_L1_CA_DLL_PLL_Tracking"), ("Tracking_1B", "Galileo_E1_DLL_PLL_VEML_Tracking"), ("Tracking_L5", "GPS_L5_DLL_PLL_Tracking"),
    ("Tracking_5X", "Galileo_E5a_DLL_PLL_Tracking"), ("Tracking_B1", "BEIDOU_B1I_DLL_PLL_Tracking"), ("Tracking_2S", "GPS_L2_M_DLL_PLL_Tracking"),
    ("Tracking_B3", "BEIDOU_B3I_DLL_PLL_Tracking"),
]
ACQ_VARIANTS = {
    "defaults": "",
    "all keys": "{r}.item_type=cshort\n{r}.dump=false\n{r}.dump_channel=2\n{r}.blocking=false\n{r}.doppler_max=7500\n{r}.coherent_integration_time_ms=2\n"
                "{r}.bit_transition_flag=false\n{r}.use_CFAR_algorithm=false\n{r}.max_dwells=3\n{r}.second_nbins=8\n{r}.second_doppler_step=62.5\n{r}.make_two_steps=true\n"
                "{r}.blocking_on_standby=true\n{r}.acquire_pilot=true\n",
    "bit transition + pfa": "{r}.bit_transition_flag=true\n{r}.max_dwells=4\n{r}.pfa=0.001\n{r}0.pfa=0.01\n",
    "resampler": "GNSS-SDR.use_acquisition_resampler=true\n{r}.coherent_integration_time_ms=1\n",
    "odd rate": "{r}.coherent_integration_time_ms=3\n{r}.acquire_iq=true\n{r}.pfa=0.0001\n",
}
TRK_VARIANTS = {
    "defaults": "",
    "all keys": "{r}.dump=false\n{r}.dump_mat=false\n{r}.high_dyn=true\n{r}.smoother_length=7\n{r}.pll_bw_hz=33.0\n{r}.dll_bw_hz=3.5\n{r}.pll_bw_narrow_hz=11.0\n"
                "{r}.dll_bw_narrow_hz=0.7\n{r}.dll_filter_order=3\n{r}.pll_filter_order=2\n{r}.enable_fll_pull_in=true\n{r}.fll_bw_hz=21.0\n{r}.pull_in_time_s=3.0\n"
                "{r}.early_late_space_chips=0.4\n{r}.early_late_space_narrow_chips=0.2\n{r}.very_early_late_space_chips=0.7\n{r}.very_early_late_space_narrow_chips=0.5\n"
                "{r}.extend_correlation_symbols=4\n{r}.track_pilot=true\n{r}.cn0_samples=30\n{r}.cn0_min=22\n{r}.max_lock_fail=40\n{r}.carrier_lock_th=0.7\n",
    "clamps": "{r}.extend_correlation_symbols=35\n{r}.dll_filter_order=7\n{r}.pll_filter_order=1\n{r}.smoother_length=0\n",
    "clamps low": "{r}.extend_correlation_symbols=0\n{r}.track_pilot=false\n{r}.dll_filter_order=0\n{r}.pll_filter_order=9\n",
}


def lib():
    if not os.path.exists(LIB):
        pytest.skip("tests/native/libb200_adapter_conf_test.so not built (needs /root/reference at build time)")
    L = C.CDLL(LIB)
    L.b200_adapter_conf_compare.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint, C.c_float, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    return L


def compare(tmp_path, kind, role, ref_impl, body, fs):
    p = tmp_path / "c.conf"
    p.write_text("[GNSS-SDR]\nGNSS-SDR.internal_fs_sps=%d\nGNSS-SDR.b200_batch_window_us=0\n" % fs + body.format(r=role))
    a, b = (C.c_double * 64)(), (C.c_double * 64)()
    n = lib().b200_adapter_conf_compare(str(p).encode(), kind, role.encode(), ref_impl.encode(), (ref_impl + "_B200").encode(), 250, 0.0045, a, b)
    assert n > 0, (role, ref_impl, n)
    return np.array(a[:n]), np.array(b[:n])


@pytest.mark.parametrize("variant", list(ACQ_VARIANTS))
@pytest.mark.parametrize("role,ref_impl", ACQ)
def test_acquisition_adapter_derives_the_reference_adapters_acq_conf(tmp_path, role, ref_impl, variant):
    fs = {"resampler": 20_000_000, "odd rate": 6_625_000}.get(variant, 8_000_000)
    if ref_impl.startswith(("Galileo_E5a", "GPS_L5", "BEIDOU_B3I")) and fs < 12_000_000:
        fs = 25_000_000
    if ref_impl.startswith("GPS_L2_M"):
        fs = min(fs, 4_000_000)  # 20 ms of code per dwell: keep the reference block's wipe-off tables small
    ref, got = compare(tmp_path, 0, role, ref_impl, ACQ_VARIANTS[variant], fs)
    assert len(ref) == len(ACQ_FIELDS)
    for name, r, g in zip(ACQ_FIELDS, ref, got):
        if name == "item_size" and "cshort" in ACQ_VARIANTS[variant] and r == 8 and g == 4:
            continue  # adapters that only know gr_complex there fall back to it; the B200 adapters take cshort / cbyte for every signal
        assert r == g, (role, ref_impl, variant, name, r, g)


@pytest.mark.parametrize("variant", list(TRK_VARIANTS))
@pytest.mark.parametrize("role,ref_impl", TRK)
def test_tracking_adapter_derives_the_reference_adapters_dll_pll_conf(tmp_path, role, ref_impl, variant):
    fs = 25_000_000 if ref_impl.startswith(("Galileo_E5a", "GPS_L5", "BEIDOU_B3I")) else 8_000_000
    ref, got = compare(tmp_path, 1, role, ref_impl, TRK_VARIANTS[variant], fs)
    assert len(ref) == len(TRK_FIELDS)
    for name, r, g in zip(TRK_FIELDS, ref, got):
        if name == "item_size" and "cshort" in TRK_VARIANTS[variant] and r == 8 and g == 4:
            continue
        assert r == g, (role, ref_impl, variant, name, r, g)
