"""ctypes access to the host-side C++ mirror (gnss-sdr-1_b200/libgnss_b200_host.so) and to the test-only build of the same
epoch loop on the oracle correlator (tests/native/libhost_oracle_loop.so)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


class TrkConf(C.Structure):
    _fields_ = [("fs_in", C.c_double), ("vector_length", C.c_int), ("pll_bw_hz", C.c_float), ("dll_bw_hz", C.c_float), ("fll_bw_hz", C.c_float),
                ("pll_filter_order", C.c_int), ("dll_filter_order", C.c_int), ("enable_fll_pull_in", C.c_int), ("enable_fll_steady_state", C.c_int),
                ("pull_in_time_s", C.c_float), ("early_late_space_chips", C.c_float), ("very_early_late_space_chips", C.c_float),
                ("cn0_samples", C.c_int), ("cn0_min", C.c_float), ("max_lock_fail", C.c_int), ("carrier_lock_th", C.c_float),
                ("track_pilot", C.c_int), ("veml", C.c_int), ("signal_carrier_freq", C.c_double), ("code_chip_rate", C.c_double),
                ("code_length_chips", C.c_int), ("code_samples_per_chip", C.c_int), ("code_period", C.c_double),
                ("pll_bw_narrow_hz", C.c_float), ("dll_bw_narrow_hz", C.c_float), ("early_late_space_narrow_chips", C.c_float),
                ("very_early_late_space_narrow_chips", C.c_float), ("extend_correlation_symbols", C.c_int), ("high_dyn", C.c_int),
                ("smoother_length", C.c_int), ("secondary", C.c_int), ("symbols_per_bit", C.c_int), ("interchange_iq", C.c_int),
                ("correlation_length_ms", C.c_int), ("prn", C.c_uint32), ("n_preamble_symbols", C.c_int), ("preamble_symbols", C.c_int32 * 200),
                ("secondary_code", C.c_char * 128), ("dump_filename", C.c_char * 256)]


class Epoch(C.Structure):
    _fields_ = [("valid", C.c_int), ("loss_of_lock", C.c_int), ("state", C.c_int), ("current_prn_length_samples", C.c_int),
                ("tracking_sample_counter", C.c_uint64), ("prompt_i", C.c_double), ("prompt_q", C.c_double), ("code_phase_samples", C.c_double),
                ("carrier_phase_rads", C.c_double), ("carrier_doppler_hz", C.c_double), ("cn0_db_hz", C.c_double), ("code_error_chips", C.c_double),
                ("carr_phase_error_hz", C.c_double), ("code_freq_chips", C.c_double), ("taps_iq", C.c_float * 12),
                ("correlation_length_ms", C.c_int), ("reserved_", C.c_int)]


GALILEO_E1_C_SECONDARY_CODE = "0011100000001010110110010"  # Galileo_E1.h:410
GPS_PREAMBLE_BITS = (1, 0, 0, 0, 1, 0, 1, 1)                # GPS_L1_CA.h:82-85


def gps_l1_conf(**kw):
    """The GPS L1 C/A adapter's defaults (gps_l1_ca_dll_pll_tracking.cc:76-186): PLL 50 Hz, DLL 2 Hz, 3rd / 2nd order, E-L 0.5 chip."""
    d = dict(fs_in=4e6, vector_length=4000, pll_bw_hz=50.0, dll_bw_hz=2.0, fll_bw_hz=35.0, pll_filter_order=3, dll_filter_order=2,
             enable_fll_pull_in=0, enable_fll_steady_state=0, pull_in_time_s=2.0, early_late_space_chips=0.5, very_early_late_space_chips=0.5,
             cn0_samples=20, cn0_min=30.0, max_lock_fail=50, carrier_lock_th=0.80, track_pilot=0, veml=0, signal_carrier_freq=1575.42e6,
             code_chip_rate=1.023e6, code_length_chips=1023, code_samples_per_chip=1, code_period=0.001,
             pll_bw_narrow_hz=20.0, dll_bw_narrow_hz=2.0, early_late_space_narrow_chips=0.5, very_early_late_space_narrow_chips=0.5,
             extend_correlation_symbols=1, high_dyn=0, smoother_length=10, secondary=0, symbols_per_bit=20, interchange_iq=0,
             correlation_length_ms=1, prn=1, n_preamble_symbols=0, secondary_code=b"", dump_filename=b"",
             preamble_symbols=[1 if b else -1 for b in GPS_PREAMBLE_BITS for _ in range(20)])
    d.update(kw)
    return _conf(d)


def _conf(d):
    d = dict(d)
    pre = d.pop("preamble_symbols", None)
    for k in ("secondary_code", "dump_filename"):
        if isinstance(d.get(k), str):
            d[k] = d[k].encode()
    c = TrkConf(**d)
    if pre is not None:
        c.n_preamble_symbols = len(pre)
        for i, v in enumerate(pre):
            c.preamble_symbols[i] = int(v)
    return c


def galileo_e1_conf(**kw):
    """The Galileo E1 adapter's defaults (galileo_e1_dll_pll_veml_tracking.cc): PLL 5 Hz, DLL 0.5 Hz, E-L 0.15, VE-VL 0.6 chip, 4 ms
    code, two table samples per chip; tracking the pilot (E1C) makes the 25-chip secondary code the synchronisation pattern."""
    d = dict(fs_in=8e6, vector_length=32000, pll_bw_hz=5.0, dll_bw_hz=0.5, fll_bw_hz=35.0, pll_filter_order=3, dll_filter_order=2,
             enable_fll_pull_in=0, enable_fll_steady_state=0, pull_in_time_s=2.0, early_late_space_chips=0.15, very_early_late_space_chips=0.6,
             cn0_samples=20, cn0_min=25.0, max_lock_fail=50, carrier_lock_th=0.85, track_pilot=1, veml=1, signal_carrier_freq=1575.42e6,
             code_chip_rate=1.023e6, code_length_chips=4092, code_samples_per_chip=2, code_period=0.004,
             pll_bw_narrow_hz=2.0, dll_bw_narrow_hz=0.25, early_late_space_narrow_chips=0.15, very_early_late_space_narrow_chips=0.6,
             extend_correlation_symbols=1, high_dyn=0, smoother_length=10, secondary=1, symbols_per_bit=1, interchange_iq=0,
             correlation_length_ms=4, prn=1, n_preamble_symbols=0, secondary_code=GALILEO_E1_C_SECONDARY_CODE, dump_filename=b"")
    d.update(kw)
    return _conf(d)


_host = None
_hol = None


def host():
    global _host
    if _host is None:
        L = C.CDLL(os.path.join(ROOT, "gnss-sdr-1_b200", "libgnss_b200_host.so"))
        for n in ("gbh_pll_cloop_two_quadrant_atan", "gbh_pll_four_quadrant_atan"):
            getattr(L, n).restype = C.c_double
            getattr(L, n).argtypes = [C.c_float, C.c_float]
        L.gbh_fll_four_quadrant_atan.restype = C.c_double
        L.gbh_fll_four_quadrant_atan.argtypes = [C.c_float] * 4 + [C.c_double] * 2
        L.gbh_dll_nc_e_minus_l_normalized.restype = C.c_double
        L.gbh_dll_nc_e_minus_l_normalized.argtypes = [C.c_float] * 4
        L.gbh_dll_nc_vemlp_normalized.restype = C.c_double
        L.gbh_dll_nc_vemlp_normalized.argtypes = [f32p] * 4
        L.gbh_cn0_svn_estimator.restype = C.c_float
        L.gbh_cn0_svn_estimator.argtypes = [f32p, C.c_int, C.c_double]
        L.gbh_carrier_lock_detector.restype = C.c_float
        L.gbh_carrier_lock_detector.argtypes = [f32p, C.c_int]
        L.gbh_loop_filter_run.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int, C.c_float, f32p, f32p, C.c_int]
        L.gbh_fll_pll_filter_run.argtypes = [C.c_float, C.c_float, C.c_int, C.c_float, f32p, f32p, C.c_float, f32p, C.c_int]
        L.gbh_trk_open.restype = C.c_void_p
        L.gbh_trk_open.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(TrkConf)]
        L.gbh_trk_start.argtypes = [C.c_void_p, f32p, C.c_void_p, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        L.gbh_trk_step.argtypes = [C.c_void_p, C.POINTER(Epoch)]
        L.gbh_trk_close.argtypes = [C.c_void_p]
        L.gbh_trk_set_channel.argtypes = [C.c_void_p, C.c_uint]
        L.gbh_matfile_write_grid.argtypes = [C.c_char_p, f32p, C.c_size_t, C.c_size_t, C.c_uint32, C.c_uint64, C.c_int32]
        L.gbh_acq_search_dump.argtypes = [C.c_void_p, C.c_void_p, C.c_int, f32p, C.c_size_t, C.c_uint64, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                          C.POINTER(C.c_uint64), C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_char_p, C.c_uint, C.c_char, C.c_char_p, C.c_uint,
                                          C.c_char_p, C.c_size_t]
        L.gbh_pcps_threshold_from_pfa.restype = C.c_float
        L.gbh_pcps_threshold_from_pfa.argtypes = [C.c_float, C.c_uint, C.c_uint, C.c_uint]
        L.gbh_acq_search.argtypes = [C.c_void_p, C.c_void_p, C.c_int, f32p, C.c_size_t, C.c_uint64, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.gbh_2nd_filter_run.argtypes = [C.c_int, C.c_float, C.c_float, f32p, f32p, C.c_int]
        L.gbh_glo_open.restype = C.c_void_p
        L.gbh_glo_open.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.gbh_glo_start.argtypes = [C.c_void_p, f32p, C.c_int, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        L.gbh_glo_step.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.gbh_glo_close.argtypes = [C.c_void_p]
        L.gbh_kf_open.restype = C.c_void_p
        L.gbh_kf_open.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.gbh_kf_start.argtypes = [C.c_void_p, f32p, C.c_double, C.c_double, C.c_uint64, C.c_uint32, C.c_uint64]
        L.gbh_kf_step.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.gbh_kf_close.argtypes = [C.c_void_p]
        L.gbh_acqv_attempt.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_int, f32p, C.c_size_t, C.c_uint64, C.c_int,
                                       C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.gbh_e5a_attempt.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, f32p, f32p, C.c_size_t, C.c_uint64, C.c_int,
                                      C.POINTER(C.c_double), C.POINTER(C.c_int), f32p]
        L.gbh_fd_attempt.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, f32p, C.c_size_t, C.c_uint64, C.POINTER(C.c_double)]
        L.gbh_caid_open.restype = C.c_void_p
        L.gbh_caid_open.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.gbh_caid_start.argtypes = [C.c_void_p, f32p, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        L.gbh_caid_set_preamble.argtypes = [C.c_void_p, C.c_double]
        L.gbh_caid_step.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.gbh_caid_close.argtypes = [C.c_void_p]
        L.gbh_acq_resampler_decimation.restype = C.c_uint
        L.gbh_acq_resampler_decimation.argtypes = [C.c_longlong, C.c_uint]
        L.gbh_attach_acq_resampler.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_uint]
        L.gbh_acq_search_resampled.argtypes = [C.c_void_p, C.c_void_p, C.c_int, f32p, C.c_size_t, C.c_uint64, C.c_int, C.c_float, C.c_uint,
                                               C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.POINTER(C.c_float)]
        L.gbh_caf_filter.argtypes = [f32p, C.c_void_p, f32p, C.c_uint, C.c_int, C.c_uint]
        _host = L
    return _host


def acq_resampler_decimation(fs_in, opt_acq_fs):
    return host().gbh_acq_resampler_decimation(fs_in, opt_acq_fs)


def attach_acq_resampler(service, src_stream, dst_stream, fs_in, decimation):
    """attach_acq_resampler of gb200_acquisition.hpp: the flowgraph's firdes taps + decimator on src_stream; returns the latency."""
    return host().gbh_attach_acq_resampler(service.h, src_stream, dst_stream, fs_in, decimation)


def acq_search_resampled(service, cfg, stream_id, code_c64, samp_count, max_runs, resampler_ratio, latency_samples):
    """B200_Pcps_Acquisition with use_automatic_resampler: returns (event, Acq_delay_samples, Acq_doppler_hz, Acq_samplestamp_samples, test_statistic)."""
    L = host()
    code = np.ascontiguousarray(code_c64, np.complex64).view(np.float32)
    d, f, st, ts = C.c_double(0), C.c_double(0), C.c_uint64(0), C.c_float(0)
    ev = L.gbh_acq_search_resampled(service.h, C.byref(cfg), stream_id, code, len(code) // 2, samp_count, max_runs, resampler_ratio, latency_samples,
                                    C.byref(d), C.byref(f), C.byref(st), C.byref(ts))
    return ev, d.value, f.value, st.value, ts.value


def caf_filter(caf_i, caf_q, caf_window_hz, doppler_step):
    """detail::caf_filter of gb200_acquisition_variants.hpp (host arithmetic only)."""
    L = host()
    vi = np.ascontiguousarray(caf_i, np.float32)
    vq = None if caf_q is None else np.ascontiguousarray(caf_q, np.float32)
    out = np.zeros(len(vi), np.float32)
    L.gbh_caf_filter(vi, None if vq is None else vq.ctypes.data_as(C.c_void_p), out, len(vi), caf_window_hz, doppler_step)
    return out


E5A_RECORD_FIELDS = ("event", "code_phase_index", "doppler_hz", "doppler_bin", "test_statistic", "input_power", "mag", "well_count",
                     "acq_delay_samples", "updated")


def e5a_caf_attempt(service, stream_id, code_i_c64, code_q_c64, samp_count, max_runs, sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms,
                    samples_per_code, threshold, max_dwells=1, bit_transition_flag=False, both_signal_components=False, caf_window_hz=0,
                    zero_padding=0, item_type=0):
    """B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf: one attempt; returns (event, records [dwells][10], CAF vector of the last dwell)."""
    L = host()
    cfg = (C.c_double * 13)(sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, max_dwells,
                            int(bit_transition_flag), int(both_signal_components), caf_window_hz, zero_padding, item_type)
    ci = np.ascontiguousarray(code_i_c64, np.complex64).view(np.float32)
    cq = np.ascontiguousarray(code_q_c64, np.complex64).view(np.float32)
    bins = len(range(-int(doppler_max), int(doppler_max) + 1, int(doppler_step)))
    caf = np.zeros(bins, np.float32)
    rec = (C.c_double * (10 * max_runs))()
    n = C.c_int(0)
    ev = L.gbh_e5a_attempt(service.h, cfg, stream_id, ci, cq, len(ci) // 2, samp_count, max_runs, rec, C.byref(n), caf)
    return ev, np.array(rec[: 10 * n.value]).reshape(n.value, 10), caf


def acq_variant_attempt(service, variant, stream_id, code_c64, samp_count, max_runs, sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms,
                        samples_per_code, threshold, max_dwells=1, tong_init_val=1, tong_max_val=2, tong_max_dwells=20, item_type=0):
    """B200_Pcps_Tong_Acquisition (variant 0) / B200_Galileo_Pcps_8ms_Acquisition (variant 1): one attempt; returns (event, records)
    with one record per dwell: event, code_phase_index, Acq_doppler_hz, test_statistic, input_power, tong_count, dwell_count, Acq_delay_samples."""
    L = host()
    cfg = (C.c_double * 12)(sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, max_dwells, tong_init_val,
                            tong_max_val, tong_max_dwells, item_type)
    code = np.ascontiguousarray(code_c64, np.complex64).view(np.float32)
    rec = (C.c_double * (8 * max_runs))()
    n = C.c_int(0)
    ev = L.gbh_acqv_attempt(service.h, variant, cfg, stream_id, code, len(code) // 2, samp_count, max_runs, rec, C.byref(n))
    return ev, np.array(rec[: 8 * n.value]).reshape(n.value, 8)


def quicksync_attempt(service, stream_id, code_c64, samp_count, max_runs, folding_factor, sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms,
                      samples_per_code, threshold, max_dwells=1, bit_transition_flag=0):
    """B200_Pcps_QuickSync_Acquisition: one attempt; returns (event, records) with one record per dwell: event, folded_index, Acq_doppler_hz,
    test_statistic, input_power, mag, well_count, Acq_delay_samples, Acq_samplestamp, best |corr|^2."""
    L = host()
    L.gbh_qs_attempt.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, f32p, C.c_size_t, C.c_uint64, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
    cfg = (C.c_double * 10)(folding_factor, sampled_ms, max_dwells, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, bit_transition_flag, threshold)
    code = np.ascontiguousarray(code_c64, np.complex64).view(np.float32)
    rec = (C.c_double * (10 * max_runs))()
    n = C.c_int(0)
    ev = L.gbh_qs_attempt(service.h, cfg, stream_id, code, len(code) // 2, samp_count, max_runs, rec, C.byref(n))
    return ev, np.array(rec[: 10 * n.value]).reshape(n.value, 10)


def oracle_loop():
    global _hol
    if _hol is None:
        L = C.CDLL(os.path.join(ROOT, "tests", "native", "libhost_oracle_loop.so"))
        L.hol_open.restype = C.c_void_p
        L.hol_open.argtypes = [f32p, C.c_uint64, C.c_uint64, C.POINTER(TrkConf)]
        L.hol_start.argtypes = [C.c_void_p, f32p, C.c_void_p, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        L.hol_step.argtypes = [C.c_void_p, C.POINTER(Epoch)]
        L.hol_close.argtypes = [C.c_void_p]
        L.hol_glo_open.restype = C.c_void_p
        L.hol_glo_open.argtypes = [f32p, C.c_uint64, C.c_uint64, C.POINTER(C.c_double)]
        L.hol_glo_start.argtypes = [C.c_void_p, f32p, C.c_int, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        L.hol_glo_step.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.hol_glo_close.argtypes = [C.c_void_p]
        L.hol_kf_open.restype = C.c_void_p
        L.hol_kf_open.argtypes = [f32p, C.c_uint64, C.c_uint64, C.POINTER(C.c_double)]
        L.hol_kf_start.argtypes = [C.c_void_p, f32p, C.c_double, C.c_double, C.c_uint64, C.c_uint32, C.c_uint64]
        L.hol_kf_step.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.hol_kf_close.argtypes = [C.c_void_p]
        L.hol_caid_open.restype = C.c_void_p
        L.hol_caid_open.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_double)]
        L.hol_caid_start.argtypes = [C.c_void_p, f32p, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        L.hol_caid_set_preamble.argtypes = [C.c_void_p, C.c_double]
        L.hol_caid_step.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.hol_caid_close.argtypes = [C.c_void_p]
        L.hol_set_channel.argtypes = [C.c_void_p, C.c_uint]
        L.hol_telemetry.argtypes = [C.c_void_p, C.c_int]
        _hol = L
    return _hol


FIELDS = ("valid", "loss_of_lock", "state", "current_prn_length_samples", "tracking_sample_counter", "prompt_i", "prompt_q", "code_phase_samples",
          "carrier_phase_rads", "carrier_doppler_hz", "cn0_db_hz", "code_error_chips", "carr_phase_error_hz", "code_freq_chips", "correlation_length_ms")


def _record(e):
    d = {k: getattr(e, k) for k in FIELDS}
    d["taps"] = np.array(e.taps_iq[:], np.float32).view(np.complex64).copy()
    return d


def run_oracle_channel(samples_c64, first_index, conf, code, acq_delay, acq_doppler, acq_stamp, n_steps, data_code=None, channel=None,
                       telemetry_fault_at=None):
    """telemetry_fault_at: step index before which the telemetry_to_trk fault message (int 1) is delivered"""
    L = oracle_loop()
    x = np.ascontiguousarray(samples_c64, np.complex64).view(np.float32)
    h = L.hol_open(x, len(x) // 2, first_index, C.byref(conf))
    if channel is not None:
        L.hol_set_channel(h, channel)
    code = np.ascontiguousarray(code, np.float32)
    dc = np.ascontiguousarray(data_code, np.float32) if data_code is not None else None
    assert L.hol_start(h, code, dc.ctypes.data_as(C.c_void_p) if dc is not None else None, acq_delay, acq_doppler, acq_stamp, first_index) == 0
    out, e = [], Epoch()
    for k in range(n_steps):
        if telemetry_fault_at is not None and k == telemetry_fault_at:
            L.hol_telemetry(h, 1)
        if L.hol_step(h, C.byref(e)) != 0:
            break
        out.append(_record(e))
    L.hol_close(h)
    return out


def run_b200_channel(service, stream_id, item_type, first_index, conf, code, acq_delay, acq_doppler, acq_stamp, n_steps, data_code=None, channel=None):
    L = host()
    h = L.gbh_trk_open(service.h, stream_id, item_type, C.byref(conf))
    assert h, "gbh_trk_open failed"
    if channel is not None:
        L.gbh_trk_set_channel(h, channel)
    code = np.ascontiguousarray(code, np.float32)
    dc = np.ascontiguousarray(data_code, np.float32) if data_code is not None else None
    assert L.gbh_trk_start(h, code, dc.ctypes.data_as(C.c_void_p) if dc is not None else None, acq_delay, acq_doppler, acq_stamp, first_index) == 0
    out, e = [], Epoch()
    for _ in range(n_steps):
        if L.gbh_trk_step(h, C.byref(e)) != 0:
            break
        out.append(_record(e))
    L.gbh_trk_close(h)
    return out


GLO_EPOCH_FIELDS = ("tracking_sample_counter", "prompt_i", "prompt_q", "carrier_doppler_hz", "code_phase_samples", "carrier_phase_rads", "cn0_db_hz",
                    "abs_e", "abs_p", "abs_l", "flags", "prn_length_samples")


def _glo_conf(fs_in, vector_length, pll_bw_hz, dll_bw_hz, early_late_space_chips, band_freq_hz=1.602e9, band_dfreq_hz=0.5625e6):
    return (C.c_double * 7)(fs_in, vector_length, pll_bw_hz, dll_bw_hz, early_late_space_chips, band_freq_hz, band_dfreq_hz)


def _glo_run(step, h, n_steps):
    out = np.zeros((n_steps, len(GLO_EPOCH_FIELDS)))
    e = (C.c_double * len(GLO_EPOCH_FIELDS))()
    for k in range(n_steps):
        assert step(h, e) == 0, "correlator refused the request at step %d" % k
        out[k] = e[:]
    return out


def run_oracle_glonass(samples_c64, first_index, code_c64, freq_channel, acq_delay, acq_doppler, acq_stamp, n_steps, **conf):
    """Glonass L1 C/A loop (gb200_glonass_tracking.hpp) on the oracle's complex-code correlator; rows of GLO_EPOCH_FIELDS."""
    L = oracle_loop()
    x = np.ascontiguousarray(samples_c64, np.complex64).view(np.float32)
    h = L.hol_glo_open(x, len(x) // 2, first_index, _glo_conf(**conf))
    assert L.hol_glo_start(h, np.ascontiguousarray(code_c64, np.complex64).view(np.float32), freq_channel, acq_delay, acq_doppler, acq_stamp, first_index) == 0
    try:
        return _glo_run(L.hol_glo_step, h, n_steps)
    finally:
        L.hol_glo_close(h)


def run_b200_glonass(service, stream_id, item_type, first_index, code_c64, freq_channel, acq_delay, acq_doppler, acq_stamp, n_steps, **conf):
    """the same loop on B200_Multicorrelator (complex codes) through the C ABI"""
    L = host()
    h = L.gbh_glo_open(service.h, stream_id, item_type, _glo_conf(**conf))
    assert h
    assert L.gbh_glo_start(h, np.ascontiguousarray(code_c64, np.complex64).view(np.float32), freq_channel, acq_delay, acq_doppler, acq_stamp, first_index) == 0
    try:
        return _glo_run(L.gbh_glo_step, h, n_steps)
    finally:
        L.gbh_glo_close(h)


KF_EPOCH_FIELDS = GLO_EPOCH_FIELDS + ("carrier_dopplerrate_hz2", "carr_phase_sigma2", "rem_carr_phase_rad")


def _kf_conf(fs_in, vector_length, dll_bw_hz, order=2, early_late_space_chips=0.5, bce_run=False, bce_ptrans=0, bce_strans=0, bce_nu=0, bce_kappa=0,
             cn0_samples=20, cn0_min=25, max_lock_fail=50, carrier_lock_th=0.85):
    return (C.c_double * 14)(order, fs_in, vector_length, dll_bw_hz, early_late_space_chips, 1.0 if bce_run else 0.0, bce_ptrans, bce_strans, bce_nu, bce_kappa,
                             cn0_samples, cn0_min, max_lock_fail, carrier_lock_th)


def _kf_run(step, h, n_steps):
    out = np.zeros((n_steps, len(KF_EPOCH_FIELDS)))
    e = (C.c_double * len(KF_EPOCH_FIELDS))()
    for k in range(n_steps):
        assert step(h, e) == 0, "correlator refused the request at step %d" % k
        out[k] = e[:]
    return out


def run_oracle_kf(samples_c64, first_index, code_f32, acq_delay, acq_doppler, acq_stamp, acq_doppler_step, n_steps, **conf):
    """Kalman GPS L1 C/A loop (gb200_kf_tracking.hpp) on the oracle's real-code correlator; rows of KF_EPOCH_FIELDS."""
    L = oracle_loop()
    x = np.ascontiguousarray(samples_c64, np.complex64).view(np.float32)
    h = L.hol_kf_open(x, len(x) // 2, first_index, _kf_conf(**conf))
    assert L.hol_kf_start(h, np.ascontiguousarray(code_f32, np.float32), acq_delay, acq_doppler, acq_stamp, acq_doppler_step, first_index) == 0
    try:
        return _kf_run(L.hol_kf_step, h, n_steps)
    finally:
        L.hol_kf_close(h)


def run_b200_kf(service, stream_id, item_type, first_index, code_f32, acq_delay, acq_doppler, acq_stamp, acq_doppler_step, n_steps, **conf):
    """the same loop on B200_Multicorrelator_Real_Codes through the C ABI"""
    L = host()
    h = L.gbh_kf_open(service.h, stream_id, item_type, _kf_conf(**conf))
    assert h
    assert L.gbh_kf_start(h, np.ascontiguousarray(code_f32, np.float32), acq_delay, acq_doppler, acq_stamp, acq_doppler_step, first_index) == 0
    try:
        return _kf_run(L.gbh_kf_step, h, n_steps)
    finally:
        L.gbh_kf_close(h)


CAID_EPOCH_FIELDS = ("tracking_sample_counter", "prompt_i", "prompt_q", "carrier_doppler_hz", "code_phase_samples", "carrier_phase_rads", "cn0_db_hz",
                     "abs_e", "abs_p", "abs_l", "flags", "correlation_length_samples", "correlation_length_ms")


def _caid_conf(fs_in, vector_length, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz=20.0, dll_bw_narrow_hz=2.0, early_late_space_chips=0.5, extend_correlation_ms=1):
    return (C.c_double * 8)(fs_in, vector_length, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_chips, extend_correlation_ms)


def _caid_run(step, set_preamble, h, n_steps, preamble):
    out = np.zeros((n_steps, len(CAID_EPOCH_FIELDS)))
    e = (C.c_double * len(CAID_EPOCH_FIELDS))()
    for k in range(n_steps):
        if preamble is not None and k == preamble[0]:
            set_preamble(h, preamble[1])
        assert step(h, e) == 0, "correlator refused the request at step %d" % k
        out[k] = e[:]
    return out


def run_oracle_caid(samples, variant, first_index, code_c64, acq_delay, acq_doppler, acq_stamp, n_steps, preamble=None, **conf):
    """Carrier-aided GPS L1 C/A loop (gb200_c_aid_tracking.hpp) on the oracle's correlators: variant 0 = _cc (complex64 samples),
    variant 1 = _sc (interleaved int16 samples).  preamble = (step index, preamble_timestamp_s) starts extended integration."""
    L = oracle_loop()
    x = np.ascontiguousarray(samples, np.int16) if variant == 1 else np.ascontiguousarray(samples, np.complex64).view(np.float32)
    h = L.hol_caid_open(x.ctypes.data_as(C.c_void_p), len(x) // 2, first_index, variant, _caid_conf(**conf))
    assert L.hol_caid_start(h, np.ascontiguousarray(code_c64, np.complex64).view(np.float32), acq_delay, acq_doppler, acq_stamp, first_index) == 0
    try:
        return _caid_run(L.hol_caid_step, L.hol_caid_set_preamble, h, n_steps, preamble)
    finally:
        L.hol_caid_close(h)


def run_b200_caid(service, stream_id, variant, first_index, code_c64, acq_delay, acq_doppler, acq_stamp, n_steps, preamble=None, **conf):
    """the same loop on B200_Multicorrelator (variant 0, 32fc stream) / B200_Multicorrelator_16sc (variant 1, 16ic stream) through the C ABI"""
    L = host()
    h = L.gbh_caid_open(service.h, stream_id, variant, _caid_conf(**conf))
    assert h
    assert L.gbh_caid_start(h, np.ascontiguousarray(code_c64, np.complex64).view(np.float32), acq_delay, acq_doppler, acq_stamp, first_index) == 0
    try:
        return _caid_run(L.gbh_caid_step, L.gbh_caid_set_preamble, h, n_steps, preamble)
    finally:
        L.gbh_caid_close(h)


FD_FIELDS = ("event", "index_time", "index_doppler", "first_peak", "second_peak", "test_statistic", "acq_delay_samples", "acq_doppler_hz", "fine_index",
             "fine_accepted")


def fine_doppler_attempt(service, stream_id, code_1ms_c64, samp_count, samples_per_ms, max_dwells, doppler_max, doppler_step, fs_in, threshold, item_type=0):
    """B200_Pcps_Acquisition_Fine_Doppler: one attempt; returns (event, dict of FD_FIELDS)."""
    L = host()
    cfg = (C.c_double * 7)(samples_per_ms, max_dwells, doppler_max, doppler_step, fs_in, threshold, item_type)
    code = np.ascontiguousarray(code_1ms_c64, np.complex64).view(np.float32)
    out = (C.c_double * len(FD_FIELDS))()
    ev = L.gbh_fd_attempt(service.h, cfg, stream_id, code, len(code) // 2, samp_count, out)
    return ev, dict(zip(FD_FIELDS, out[:]))
