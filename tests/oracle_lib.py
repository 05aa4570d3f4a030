"""ctypes doors into the CPU oracle (TEST INFRASTRUCTURE ONLY).

`oracle()`  -> oracle/libgnss_oracle.so  : plain-C restatement (oracle/gnss_oracle.c)
`ref()`     -> oracle/_ref/libgnss_ref.so: the reference's own sources compiled from /root/reference
               (prebuilt file on the GPU box; None when it was never built)

Nothing in the product package imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


class GoAcqConf(C.Structure):
    _fields_ = [
        ("sampled_ms", C.c_uint32), ("ms_per_code", C.c_uint32), ("samples_per_chip", C.c_uint32),
        ("max_dwells", C.c_uint32), ("doppler_max", C.c_uint32), ("doppler_step", C.c_uint32),
        ("num_doppler_bins_step2", C.c_uint32), ("doppler_step2", C.c_float), ("fs_in", C.c_int64),
        ("samples_per_ms", C.c_float), ("samples_per_code", C.c_float), ("bit_transition_flag", C.c_int),
        ("use_CFAR_algorithm_flag", C.c_int), ("make_2_steps", C.c_int), ("cshort", C.c_int),
        ("threshold", C.c_float), ("freq_offset_hz", C.c_int64), ("fft_f64", C.c_int),
        ("analytic_wipeoff", C.c_int), ("keep_grid", C.c_int), ("accumulator_lanes", C.c_int),
    ]


class GoAcqResult(C.Structure):
    _fields_ = [
        ("code_phase_index", C.c_uint32), ("doppler_hz", C.c_int32), ("doppler_bin", C.c_uint32),
        ("peak_mag", C.c_float), ("input_power", C.c_float), ("test_statistic", C.c_float),
        ("decision", C.c_int), ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double),
        ("acq_samplestamp", C.c_uint64),
    ]


_oracle = None
_ref = None


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "all"])


def oracle():
    global _oracle
    if _oracle is not None:
        return _oracle
    path = os.path.join(ORACLE_DIR, "libgnss_oracle.so")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(ORACLE_DIR, "gnss_oracle.c")):
        build_oracle()
    L = C.CDLL(path)
    L.go_resampler_real.argtypes = [f32p, f32p, C.c_float, C.c_float, f32p, C.c_int, C.c_int, C.c_int]
    L.go_resampler_real_hd.argtypes = [f32p, f32p, C.c_float, C.c_float, C.c_float, f32p, C.c_int, C.c_int, C.c_int]
    L.go_rotator_dot_prod_real.argtypes = [f32p, f32p, C.c_float, C.c_float, f32p, f32p, C.c_int, C.c_int]
    L.go_multicorrelator_real.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int] + [C.c_float] * 6 + [C.c_int, f32p]
    L.go_multicorrelator_cplx.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int] + [C.c_float] * 4 + [f32p]
    L.go_multicorrelator_real_f64.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int] + [C.c_float] * 4 + [f64p]
    L.go_sincos.argtypes = [f32p, C.c_float, C.c_float, C.c_int]
    L.go_index_max.argtypes = [f32p, C.c_uint32]
    L.go_index_max.restype = C.c_uint32
    L.go_16ic_convert_32fc.argtypes = [f32p, i16p, C.c_uint32]
    L.go_fft_f32.argtypes = [f32p, C.c_int, C.c_int]
    L.go_fft_f64.argtypes = [f64p, C.c_int, C.c_int]
    L.go_gps_l1_ca_code_int.argtypes = [i32p, C.c_int32, C.c_uint32]
    L.go_gps_l1_ca_code_float.argtypes = [f32p, C.c_int32, C.c_uint32]
    L.go_gps_l1_ca_code_sampled.argtypes = [f32p, C.c_uint32, C.c_int32, C.c_uint32]
    L.go_code_resampler_f32.argtypes = [f32p, f32p, C.c_float, C.c_float, C.c_uint32, C.c_uint32]
    L.go_acq_create.argtypes = [C.POINTER(GoAcqConf)]
    L.go_acq_create.restype = C.c_void_p
    L.go_acq_destroy.argtypes = [C.c_void_p]
    for fn in (L.go_acq_fft_size, L.go_acq_consumed_samples, L.go_acq_num_doppler_bins):
        fn.argtypes = [C.c_void_p]
        fn.restype = C.c_uint32
    L.go_acq_set_local_code.argtypes = [C.c_void_p, f32p]
    L.go_acq_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(GoAcqResult)]
    L.go_acq_grid.argtypes = [C.c_void_p, C.c_int]
    L.go_acq_grid.restype = C.POINTER(C.c_float)
    L.go_acqv_create.argtypes = [C.POINTER(GoAcqvConf)]
    L.go_acqv_create.restype = C.c_void_p
    L.go_acqv_destroy.argtypes = [C.c_void_p]
    L.go_acqv_num_doppler_bins.argtypes = [C.c_void_p]
    L.go_acqv_num_doppler_bins.restype = C.c_uint32
    L.go_acqv_phase_step.argtypes = [C.c_void_p, C.c_uint32]
    L.go_acqv_phase_step.restype = C.c_float
    L.go_acqv_set_local_code.argtypes = [C.c_void_p, f32p]
    L.go_acqv_restart.argtypes = [C.c_void_p]
    L.go_acqv_run.argtypes = [C.c_void_p, f32p, C.POINTER(GoAcqvResult)]
    L.go_e5a_create.argtypes = [C.POINTER(GoE5aConf)]
    L.go_e5a_create.restype = C.c_void_p
    L.go_e5a_destroy.argtypes = [C.c_void_p]
    L.go_e5a_num_doppler_bins.argtypes = [C.c_void_p]
    L.go_e5a_num_doppler_bins.restype = C.c_uint32
    L.go_e5a_phase_step.argtypes = [C.c_void_p, C.c_uint32]
    L.go_e5a_phase_step.restype = C.c_float
    L.go_e5a_set_local_code.argtypes = [C.c_void_p, f32p, f32p]
    L.go_e5a_restart.argtypes = [C.c_void_p]
    L.go_e5a_run.argtypes = [C.c_void_p, f32p, C.POINTER(GoE5aResult)]
    L.go_fd_create.argtypes = [C.POINTER(GoFdConf)]
    L.go_fd_create.restype = C.c_void_p
    L.go_fd_destroy.argtypes = [C.c_void_p]
    L.go_fd_num_doppler_points.argtypes = [C.c_void_p]
    L.go_fd_phase_step.argtypes = [C.c_void_p, C.c_int]
    L.go_fd_phase_step.restype = C.c_float
    L.go_fd_set_local_code.argtypes = [C.c_void_p, f32p]
    L.go_fd_attempt.argtypes = [C.c_void_p, f32p, f32p, C.POINTER(GoFdResult)]
    L.go_fd_replica.argtypes = [f32p, C.c_uint32, C.c_int, f32p]
    L.go_firdes_low_pass.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, f32p, C.c_int]
    L.go_fir_decimate.argtypes = [f32p, C.c_int, f32p, C.c_int, C.c_int, f32p]
    L.go_acq_resampler_decimation.argtypes = [C.c_int64, C.c_uint32]
    L.go_acq_resampler_decimation.restype = C.c_uint32
    L.go_e5a_caf_filter.argtypes = [f32p, f32p, f32p, C.c_uint32, C.c_int, C.c_uint32, C.c_int]
    L.go_e5a_caf_vector.argtypes = [C.c_void_p]
    L.go_e5a_caf_vector.restype = C.POINTER(C.c_float)
    _oracle = L
    return L


def ref():
    """The reference compiled from its own sources, or None if oracle/_ref was never built."""
    global _ref
    if _ref is not None:
        return _ref
    path = os.path.join(ORACLE_DIR, "_ref", "libgnss_ref.so")
    if not os.path.exists(path):
        if os.path.isdir("/root/reference/src"):
            subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref"])
        else:
            return None
    L = C.CDLL(path)
    L.gbref_set_arch.argtypes = [C.c_int]
    L.gbref_multicorrelator_real.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int] + [C.c_float] * 6 + [C.c_int, f32p]
    L.gbref_multicorrelator_cplx.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int] + [C.c_float] * 4 + [f32p]
    L.gbref_multicorrelator_16sc.argtypes = [i16p, C.c_int, i16p, C.c_int, f32p, C.c_int] + [C.c_float] * 4 + [i16p]
    L.gbref_resampler_real.argtypes = [f32p, f32p, C.c_float, C.c_float, f32p, C.c_int, C.c_int, C.c_int, C.c_int]
    L.gbref_rotator_dot_prod_real.argtypes = [f32p, f32p, C.c_float, C.c_float, f32p, f32p, C.c_int, C.c_int, C.c_int]
    L.gbref_sincos.argtypes = [f32p, C.c_float, C.c_float, C.c_int, C.c_int]
    L.gbref_index_max.argtypes = [f32p, C.c_uint]
    L.gbref_index_max.restype = C.c_uint
    L.gbref_16ic_convert_32fc.argtypes = [f32p, i16p, C.c_uint]
    L.gbref_gps_l1_ca_code_float.argtypes = [f32p, C.c_int, C.c_uint]
    L.gbref_gps_l1_ca_code_sampled.argtypes = [f32p, C.c_uint, C.c_int, C.c_uint]
    L.gbref_galileo_e1_code_sinboc11_float.argtypes = [f32p, C.c_char_p, C.c_uint]
    L.gbref_galileo_e1_code_sampled.argtypes = [f32p, C.c_char_p, C.c_int, C.c_uint, C.c_int, C.c_uint, C.c_int]
    L.gbref_gps_l5i_code_float.argtypes = [f32p, C.c_uint]
    L.gbref_gps_l5q_code_float.argtypes = [f32p, C.c_uint]
    L.gbref_gps_l5i_code_sampled.argtypes = [f32p, C.c_uint, C.c_int]
    L.gbref_gps_l2c_m_code_float.argtypes = [f32p, C.c_uint]
    L.gbref_galileo_e5a_code_primary.argtypes = [f32p, C.c_int, C.c_char_p]
    L.gbref_galileo_e5a_code_sampled.argtypes = [f32p, C.c_char_p, C.c_uint, C.c_int, C.c_uint]
    L.gbref_beidou_b1i_code_float.argtypes = [f32p, C.c_int, C.c_uint]
    L.gbref_beidou_b1i_code_sampled.argtypes = [f32p, C.c_uint, C.c_int, C.c_uint]
    L.gbref_glonass_l1_ca_code.argtypes = [f32p, C.c_uint]
    L.gbref_glonass_l1_ca_code_sampled.argtypes = [f32p, C.c_int, C.c_uint]
    L.gbref_fll_four_quadrant_atan.argtypes = [C.c_float] * 4 + [C.c_double] * 2
    L.gbref_fll_four_quadrant_atan.restype = C.c_double
    for fn in (L.gbref_pll_four_quadrant_atan, L.gbref_pll_cloop_two_quadrant_atan):
        fn.argtypes = [C.c_float] * 2
        fn.restype = C.c_double
    L.gbref_dll_nc_e_minus_l_normalized.argtypes = [C.c_float] * 4
    L.gbref_dll_nc_e_minus_l_normalized.restype = C.c_double
    L.gbref_dll_nc_vemlp_normalized.argtypes = [f32p] * 4
    L.gbref_dll_nc_vemlp_normalized.restype = C.c_double
    L.gbref_cn0_svn_estimator.argtypes = [f32p, C.c_int, C.c_double]
    L.gbref_cn0_svn_estimator.restype = C.c_float
    L.gbref_carrier_lock_detector.argtypes = [f32p, C.c_int]
    L.gbref_carrier_lock_detector.restype = C.c_float
    L.gbref_tracking_loop_filter_run.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int, C.c_float, f32p, f32p, C.c_int]
    if hasattr(L, "gbref_2nd_filter_run"):
        L.gbref_2nd_filter_run.argtypes = [C.c_int, C.c_float, C.c_float, f32p, f32p, C.c_int]
    L.gbref_fll_pll_filter_run.argtypes = [C.c_float, C.c_float, C.c_int, C.c_float, f32p, f32p, C.c_float, f32p, C.c_int]
    L.gbref_bench_multicorrelator_real.argtypes = [f32p, C.c_int, f32p, C.c_int, f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.gbref_bench_multicorrelator_real.restype = C.c_double
    _ref = L
    return L


# ---------------------------------------------------------------- convenience wrappers
def multicorrelator_real(lib_kind, x, code, shifts, rem_carr, phase_step, rem_code, code_step,
                         phase_rate=0.0, code_rate=0.0, high_dyn=False):
    """x: complex64 array; returns complex64[T].  lib_kind: 'oracle' | 'ref_generic' | 'ref_avx'."""
    x = np.ascontiguousarray(x, np.complex64).view(np.float32)
    code = np.ascontiguousarray(code, np.float32)
    shifts = np.ascontiguousarray(shifts, np.float32)
    out = np.zeros(2 * len(shifts), np.float32)
    args = (x, len(x) // 2, code, len(code), shifts, len(shifts), rem_carr, phase_step, phase_rate,
            rem_code, code_step, code_rate, int(high_dyn), out)
    if lib_kind == "oracle":
        oracle().go_multicorrelator_real(*args)
    else:
        r = ref()
        r.gbref_set_arch(1 if lib_kind == "ref_avx" else 0)
        r.gbref_multicorrelator_real(*args)
        r.gbref_set_arch(0)
    return out.view(np.complex64)


def acq_conf(**kw):
    c = GoAcqConf()
    d = dict(sampled_ms=1, ms_per_code=1, samples_per_chip=4, max_dwells=1, doppler_max=5000, doppler_step=250,
             num_doppler_bins_step2=4, doppler_step2=125.0, fs_in=4000000, samples_per_ms=4000.0,
             samples_per_code=4000.0, bit_transition_flag=0, use_CFAR_algorithm_flag=1, make_2_steps=0,
             cshort=0, threshold=0.0, freq_offset_hz=0, fft_f64=0, analytic_wipeoff=0, keep_grid=0, accumulator_lanes=0)
    d.update(kw)
    for k, v in d.items():
        setattr(c, k, v)
    return c


class GoAcqvConf(C.Structure):
    _fields_ = [("variant", C.c_int), ("fft_size", C.c_uint32), ("samples_per_code", C.c_uint32), ("doppler_max", C.c_uint32),
                ("doppler_step", C.c_uint32), ("fs_in", C.c_int64), ("threshold", C.c_float), ("max_dwells", C.c_uint32),
                ("tong_init_val", C.c_uint32), ("tong_max_val", C.c_uint32), ("tong_max_dwells", C.c_uint32)]


class GoAcqvResult(C.Structure):
    _fields_ = [("code_phase_index", C.c_uint32), ("doppler_hz", C.c_int32), ("doppler_bin", C.c_uint32), ("mag", C.c_float),
                ("input_power", C.c_float), ("test_statistic", C.c_float), ("decision", C.c_int), ("tong_count", C.c_int32),
                ("dwell_count", C.c_uint32), ("acq_delay_samples", C.c_double)]


ACQV_TONG, ACQV_8MS = 0, 1


class GoFdConf(C.Structure):
    _fields_ = [("samples_per_ms", C.c_uint32), ("max_dwells", C.c_uint32), ("doppler_max", C.c_uint32), ("doppler_step", C.c_uint32),
                ("fs_in", C.c_int64), ("threshold", C.c_float)]


class GoFdResult(C.Structure):
    _fields_ = [("index_time", C.c_uint32), ("index_doppler", C.c_uint32), ("first_peak", C.c_float), ("second_peak", C.c_float),
                ("test_statistic", C.c_float), ("positive", C.c_int), ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double),
                ("fine_index", C.c_uint32), ("fine_accepted", C.c_int)]


class OracleFineDoppler:
    """pcps_acquisition_fine_doppler_cc restatement (oracle/gnss_oracle.c go_fd_*): one attempt per call."""

    def __init__(self, samples_per_ms, max_dwells, doppler_max, doppler_step, fs_in, threshold):
        self.L = oracle()
        self.conf = GoFdConf(samples_per_ms, max_dwells, doppler_max, doppler_step, fs_in, threshold)
        self.h = self.L.go_fd_create(C.byref(self.conf))
        self.n = samples_per_ms
        self.num_points = self.L.go_fd_num_doppler_points(self.h)

    def phase_steps(self):
        return np.array([self.L.go_fd_phase_step(self.h, b) for b in range(self.num_points)], np.float32)

    def attempt(self, samples_c64, code_1ms_c64):
        code = np.ascontiguousarray(code_1ms_c64, np.complex64).view(np.float32)
        assert len(code) == 2 * self.n
        x = np.ascontiguousarray(samples_c64, np.complex64).view(np.float32)
        assert len(x) >= 2 * self.n * max(10, self.conf.max_dwells)
        self.L.go_fd_set_local_code(self.h, code)
        r = GoFdResult()
        self.L.go_fd_attempt(self.h, x, code, C.byref(r))
        return r

    def close(self):
        if self.h:
            self.L.go_fd_destroy(self.h)
            self.h = None


def fd_replica(code_1ms_c64, shift_index):
    code = np.ascontiguousarray(code_1ms_c64, np.complex64).view(np.float32)
    n = len(code) // 2
    out = np.zeros(20 * n, np.float32)
    oracle().go_fd_replica(code, n, shift_index, out)
    return out.view(np.complex64)


class GoE5aConf(C.Structure):
    _fields_ = [("sampled_ms", C.c_uint32), ("samples_per_ms", C.c_uint32), ("samples_per_code", C.c_uint32), ("max_dwells", C.c_uint32),
                ("doppler_max", C.c_uint32), ("doppler_step", C.c_uint32), ("fs_in", C.c_int64), ("bit_transition_flag", C.c_int),
                ("both_signal_components", C.c_int), ("caf_window_hz", C.c_int), ("zero_padding", C.c_int), ("threshold", C.c_float)]


class GoE5aResult(C.Structure):
    _fields_ = [("code_phase_index", C.c_uint32), ("doppler_bin", C.c_uint32), ("doppler_hz", C.c_int32), ("mag", C.c_float),
                ("input_power", C.c_float), ("test_statistic", C.c_float), ("decision", C.c_int), ("well_count", C.c_uint32),
                ("acq_delay_samples", C.c_double), ("updated", C.c_int)]


def firdes_low_pass(gain, fs, cutoff, transition):
    """oracle restatement of gr::filter::firdes::low_pass(.., WIN_HAMMING)"""
    L = oracle()
    n = -L.go_firdes_low_pass(gain, fs, cutoff, transition, np.zeros(1, np.float32), 0)
    taps = np.zeros(n, np.float32)
    assert L.go_firdes_low_pass(gain, fs, cutoff, transition, taps, n) == n
    return taps


def fir_decimate(x_c64, taps, decimation):
    """oracle restatement of gr::filter::fir_filter_ccf(decimation, taps) over a stream that starts at x[0]"""
    x = np.ascontiguousarray(x_c64, np.complex64).view(np.float32)
    t = np.ascontiguousarray(taps, np.float32)
    n_in = len(x) // 2
    out = np.zeros(2 * ((n_in - 1) // decimation + 1), np.float32)
    n = oracle().go_fir_decimate(x, n_in, t, len(t), decimation, out)
    return out[: 2 * n].view(np.complex64)


def e5a_caf_filter(caf_i, caf_q, caf_window_hz, doppler_step):
    vi = np.ascontiguousarray(caf_i, np.float32)
    vq = np.ascontiguousarray(caf_q if caf_q is not None else np.zeros(len(vi)), np.float32)
    out = np.zeros(len(vi), np.float32)
    oracle().go_e5a_caf_filter(vi, vq, out, len(vi), caf_window_hz, doppler_step, int(caf_q is not None))
    return out


class OracleE5aCaf:
    """galileo_e5a_noncoherentIQ_acquisition_caf_cc restatement (oracle/gnss_oracle.c go_e5a_*)."""

    def __init__(self, sampled_ms, samples_per_ms, samples_per_code, max_dwells, doppler_max, doppler_step, fs_in, threshold,
                 bit_transition_flag=False, both_signal_components=False, caf_window_hz=0, zero_padding=0):
        self.L = oracle()
        self.conf = GoE5aConf(sampled_ms, samples_per_ms, samples_per_code, max_dwells, doppler_max, doppler_step, fs_in,
                              int(bit_transition_flag), int(both_signal_components), caf_window_hz, zero_padding, threshold)
        self.h = self.L.go_e5a_create(C.byref(self.conf))
        self.fft_size = sampled_ms * samples_per_ms
        self.num_bins = self.L.go_e5a_num_doppler_bins(self.h)

    def phase_steps(self):
        return np.array([self.L.go_e5a_phase_step(self.h, b) for b in range(self.num_bins)], np.float32)

    def set_local_code(self, code_i_c64, code_q_c64):
        ci = np.ascontiguousarray(code_i_c64, np.complex64).view(np.float32)
        cq = np.ascontiguousarray(code_q_c64, np.complex64).view(np.float32)
        assert len(ci) == 2 * self.fft_size and len(cq) == 2 * self.fft_size
        self.L.go_e5a_set_local_code(self.h, ci, cq)

    def restart(self):
        self.L.go_e5a_restart(self.h)

    def run(self, samples_c64):
        x = np.ascontiguousarray(samples_c64[: self.fft_size], np.complex64).view(np.float32)
        assert len(x) == 2 * self.fft_size
        r = GoE5aResult()
        self.L.go_e5a_run(self.h, x, C.byref(r))
        return r

    def caf_vector(self):
        p = self.L.go_e5a_caf_vector(self.h)
        return np.array([p[i] for i in range(self.num_bins)], np.float32)

    def close(self):
        if self.h:
            self.L.go_e5a_destroy(self.h)
            self.h = None


class OracleAcqVariant:
    """pcps_tong_acquisition_cc / galileo_pcps_8ms_acquisition_cc restatement (oracle/gnss_oracle.c go_acqv_*)."""

    def __init__(self, variant, fft_size, samples_per_code, doppler_max, doppler_step, fs_in, threshold, max_dwells=1, tong_init_val=1,
                 tong_max_val=2, tong_max_dwells=20):
        self.L = oracle()
        self.conf = GoAcqvConf(variant, fft_size, samples_per_code, doppler_max, doppler_step, fs_in, threshold, max_dwells, tong_init_val,
                               tong_max_val, tong_max_dwells)
        self.h = self.L.go_acqv_create(C.byref(self.conf))
        self.fft_size = fft_size
        self.num_bins = self.L.go_acqv_num_doppler_bins(self.h)

    def phase_steps(self):
        return np.array([self.L.go_acqv_phase_step(self.h, b) for b in range(self.num_bins)], np.float32)

    def set_local_code(self, code_c64):
        code = np.ascontiguousarray(code_c64, np.complex64).view(np.float32)
        assert len(code) == 2 * self.fft_size
        self.L.go_acqv_set_local_code(self.h, code)

    def restart(self):
        self.L.go_acqv_restart(self.h)

    def run(self, samples_c64):
        x = np.ascontiguousarray(samples_c64[: self.fft_size], np.complex64).view(np.float32)
        r = GoAcqvResult()
        self.L.go_acqv_run(self.h, x, C.byref(r))
        return r

    def close(self):
        if self.h:
            self.L.go_acqv_destroy(self.h)
            self.h = None


class GoQsConf(C.Structure):
    _fields_ = [("folding_factor", C.c_uint32), ("sampled_ms", C.c_uint32), ("max_dwells", C.c_uint32), ("doppler_max", C.c_uint32),
                ("doppler_step", C.c_uint32), ("fs_in", C.c_int64), ("samples_per_ms", C.c_int32), ("samples_per_code", C.c_int32),
                ("bit_transition_flag", C.c_int), ("threshold", C.c_float)]


class GoQsResult(C.Structure):
    _fields_ = [("well_count", C.c_uint32), ("folded_index", C.c_uint32), ("doppler_bin", C.c_uint32), ("best_candidate", C.c_uint32),
                ("improved", C.c_int), ("mag", C.c_float), ("input_power", C.c_float), ("test_statistic", C.c_float), ("state", C.c_int),
                ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp", C.c_uint64), ("corr_output", C.c_float * 16)]


class OracleQuickSync:
    """pcps_quicksync_acquisition_cc restatement (oracle/gnss_oracle.c go_qs_*)."""

    def __init__(self, folding_factor, sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, max_dwells=1,
                 bit_transition_flag=0):
        self.L = oracle()
        L = self.L
        L.go_qs_create.argtypes = [C.POINTER(GoQsConf)]
        L.go_qs_create.restype = C.c_void_p
        L.go_qs_destroy.argtypes = [C.c_void_p]
        L.go_qs_fft_size.argtypes = [C.c_void_p]
        L.go_qs_fft_size.restype = C.c_uint32
        L.go_qs_num_doppler_bins.argtypes = [C.c_void_p]
        L.go_qs_num_doppler_bins.restype = C.c_uint32
        L.go_qs_phase_step.argtypes = [C.c_void_p, C.c_uint32]
        L.go_qs_phase_step.restype = C.c_float
        L.go_qs_set_local_code.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.go_qs_restart.argtypes = [C.c_void_p]
        L.go_qs_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(GoQsResult)]
        self.conf = GoQsConf(folding_factor, sampled_ms, max_dwells, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, bit_transition_flag, threshold)
        self.h = L.go_qs_create(C.byref(self.conf))
        self.fft_size = L.go_qs_fft_size(self.h)
        self.num_bins = L.go_qs_num_doppler_bins(self.h)
        self.vector_items = sampled_ms * samples_per_ms
        self.samples_per_code = samples_per_code

    def phase_steps(self):
        return np.array([self.L.go_qs_phase_step(self.h, b) for b in range(self.num_bins)], np.float32)

    def set_local_code(self, code_c64):
        code = np.ascontiguousarray(code_c64, np.complex64)
        assert len(code) >= self.samples_per_code
        folded = np.zeros(self.fft_size, np.complex64)
        self.L.go_qs_set_local_code(self.h, code.ctypes.data_as(C.c_void_p), folded.ctypes.data_as(C.c_void_p))
        return folded

    def restart(self):
        self.L.go_qs_restart(self.h)

    def run(self, vector_c64, sample_counter_after):
        x = np.ascontiguousarray(vector_c64[: self.vector_items], np.complex64)
        assert len(x) == self.vector_items
        r = GoQsResult()
        self.L.go_qs_run(self.h, x.ctypes.data_as(C.c_void_p), sample_counter_after, C.byref(r))
        return r

    def close(self):
        if self.h:
            self.L.go_qs_destroy(self.h)
            self.h = None


class OracleAcq:
    """pcps_acquisition restatement as an object (oracle/gnss_oracle.c go_acq_*)."""

    def __init__(self, **kw):
        self.L = oracle()
        self.conf = acq_conf(**kw)
        self.h = self.L.go_acq_create(C.byref(self.conf))
        self.fft_size = self.L.go_acq_fft_size(self.h)
        self.consumed = self.L.go_acq_consumed_samples(self.h)
        self.num_bins = self.L.go_acq_num_doppler_bins(self.h)

    def set_local_code(self, code_c64):
        code = np.ascontiguousarray(code_c64, np.complex64).view(np.float32)
        need = self.fft_size // 2 if self.conf.bit_transition_flag else self.consumed
        assert len(code) // 2 >= need, (len(code) // 2, need)
        self.L.go_acq_set_local_code(self.h, code)

    def run(self, samples, samp_count=0):
        if self.conf.cshort:
            s = np.ascontiguousarray(samples, np.int16)
            assert s.size >= 2 * self.consumed
        else:
            s = np.ascontiguousarray(samples, np.complex64).view(np.float32)
            assert s.size >= 2 * self.consumed
        r = GoAcqResult()
        self.L.go_acq_run(self.h, s.ctypes.data_as(C.c_void_p), samp_count, C.byref(r))
        return r

    def grid(self):
        eff = self.fft_size // 2 if self.conf.bit_transition_flag else self.fft_size
        return np.stack([np.ctypeslib.as_array(self.L.go_acq_grid(self.h, b), shape=(self.fft_size,))[:eff].copy()
                         for b in range(self.num_bins)])

    def close(self):
        if self.h:
            self.L.go_acq_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------- the reference's GNU Radio blocks themselves (oracle/ref_blocks.cc)
class GbRefAcqConf(C.Structure):
    _fields_ = [("sampled_ms", C.c_uint32), ("ms_per_code", C.c_uint32), ("samples_per_chip", C.c_uint32), ("max_dwells", C.c_uint32),
                ("doppler_max", C.c_uint32), ("doppler_step", C.c_uint32), ("num_doppler_bins_step2", C.c_uint32), ("doppler_step2", C.c_float),
                ("fs_in", C.c_int64), ("samples_per_ms", C.c_float), ("samples_per_code", C.c_float), ("bit_transition_flag", C.c_int),
                ("use_CFAR_algorithm_flag", C.c_int), ("make_2_steps", C.c_int), ("cshort", C.c_int), ("threshold", C.c_float),
                ("system", C.c_char), ("signal", C.c_char * 3), ("prn", C.c_uint32), ("use_fsm", C.c_int)]


class GbRefAcqState(C.Structure):
    _fields_ = [("state", C.c_int32), ("active", C.c_int32), ("step_two", C.c_int32), ("positive_acq", C.c_int32), ("dwell_counter", C.c_uint32),
                ("fft_size", C.c_uint32), ("consumed_samples", C.c_uint32), ("num_doppler_bins", C.c_uint32), ("threshold", C.c_float), ("mag", C.c_float),
                ("input_power", C.c_float), ("test_statistics", C.c_float), ("doppler_center_step_two", C.c_float), ("sample_counter", C.c_uint64),
                ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("acq_doppler_step", C.c_uint32),
                ("fsm_valid_acquisitions", C.c_int32), ("old_freq", C.c_int64)]


class GbRefTrkConf(C.Structure):
    _fields_ = [("fs_in", C.c_double), ("vector_length", C.c_uint32), ("system", C.c_char), ("signal", C.c_char * 3), ("pll_filter_order", C.c_int),
                ("dll_filter_order", C.c_int), ("fll_filter_order", C.c_int), ("enable_fll_pull_in", C.c_int), ("enable_fll_steady_state", C.c_int),
                ("pull_in_time_s", C.c_uint32), ("fll_bw_hz", C.c_float), ("pll_bw_hz", C.c_float), ("dll_bw_hz", C.c_float), ("pll_bw_narrow_hz", C.c_float),
                ("dll_bw_narrow_hz", C.c_float), ("early_late_space_chips", C.c_float), ("very_early_late_space_chips", C.c_float),
                ("early_late_space_narrow_chips", C.c_float), ("very_early_late_space_narrow_chips", C.c_float), ("extend_correlation_symbols", C.c_int32),
                ("high_dyn", C.c_int), ("smoother_length", C.c_uint32), ("cn0_samples", C.c_int32), ("cn0_min", C.c_int32), ("max_lock_fail", C.c_int32),
                ("carrier_lock_th", C.c_double), ("track_pilot", C.c_int), ("dump", C.c_int), ("dump_filename", C.c_char * 256)]


class GbRefTrkEpoch(C.Structure):
    _fields_ = [("tracking_sample_counter", C.c_uint64), ("first_sample", C.c_uint64), ("n_samples", C.c_int32), ("state_after", C.c_int32),
                ("prompt_i", C.c_double), ("prompt_q", C.c_double), ("carrier_phase_rads", C.c_double), ("carrier_doppler_hz", C.c_double),
                ("code_phase_samples", C.c_double), ("cn0_db_hz", C.c_double), ("code_freq_chips", C.c_double), ("rem_code_phase_chips", C.c_double),
                ("acc_carrier_phase_rad", C.c_double), ("carrier_lock_test", C.c_double), ("rem_carr_phase_rad", C.c_float), ("correlation_length_ms", C.c_int32),
                ("cloop", C.c_int32), ("pull_in_transitory", C.c_int32), ("current_symbol", C.c_int32), ("lock_fail_counter", C.c_int32),
                ("corr", C.c_float * 10), ("p_accu", C.c_float * 2)]


def ref_blocks():
    """oracle/_ref with the block harness bound, or None (no oracle/_ref, or one built before the blocks were added)."""
    L = ref()
    if L is None or not hasattr(L, "gbref_acq_block_create"):
        return None
    if getattr(L, "_blocks_bound", False):
        return L
    L.gbref_acq_block_create.argtypes = [C.POINTER(GbRefAcqConf)]
    L.gbref_acq_block_create.restype = C.c_void_p
    L.gbref_acq_block_destroy.argtypes = [C.c_void_p]
    L.gbref_acq_block_set_local_code.argtypes = [C.c_void_p, f32p]
    L.gbref_acq_block_set_state.argtypes = [C.c_void_p, C.c_int]
    L.gbref_acq_block_set_active.argtypes = [C.c_void_p, C.c_int]
    L.gbref_acq_block_set_threshold.argtypes = [C.c_void_p, C.c_float]
    L.gbref_acq_block_set_sample_counter.argtypes = [C.c_void_p, C.c_uint64]
    L.gbref_set_accumulator_lanes.argtypes = [C.c_int]
    L.gbref_acq_block_work.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.gbref_acq_block_get_state.argtypes = [C.c_void_p, C.POINTER(GbRefAcqState)]
    L.gbref_acq_block_grid_row.argtypes = [C.c_void_p, C.c_uint32]
    L.gbref_acq_block_grid_row.restype = C.POINTER(C.c_float)
    L.gbref_trk_block_create.argtypes = [C.POINTER(GbRefTrkConf)]
    L.gbref_trk_block_create.restype = C.c_void_p
    L.gbref_trk_block_destroy.argtypes = [C.c_void_p]
    L.gbref_trk_block_start.argtypes = [C.c_void_p, C.c_uint32, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
    L.gbref_trk_block_stop.argtypes = [C.c_void_p]
    L.gbref_trk_block_telemetry_fault.argtypes = [C.c_void_p]
    L.gbref_trk_block_state.argtypes = [C.c_void_p]
    L.gbref_trk_block_position.argtypes = [C.c_void_p]
    L.gbref_trk_block_position.restype = C.c_longlong
    L.gbref_trk_block_run.argtypes = [C.c_void_p, f32p, C.c_longlong, C.POINTER(GbRefTrkEpoch), C.c_int, i32p, C.c_int, C.POINTER(C.c_int)]
    L.gbref_read_tracking_dump.argtypes = [C.c_char_p, f64p, C.c_longlong]
    L.gbref_read_tracking_dump.restype = C.c_longlong
    L._blocks_bound = True
    return L


DUMP_READER_FIELDS = ("abs_VE", "abs_E", "abs_P", "abs_L", "abs_VL", "prompt_I", "prompt_Q", "PRN_start_sample_count", "acc_carrier_phase_rad",
                      "carrier_doppler_hz", "carrier_doppler_rate_hz_s", "code_freq_chips", "code_freq_rate_chips", "carr_error_hz", "carr_error_filt_hz",
                      "code_error_chips", "code_error_filt_chips", "CN0_SNV_dB_Hz", "carrier_lock_test", "aux1", "aux2")


def read_tracking_dump(path, max_epochs=100000):
    """A tracking .dat dump read by the reference's own Tracking_Dump_Reader; returns (num_epochs, array [n, 21])."""
    L = ref_blocks()
    out = np.zeros((max_epochs, 21), np.float64)
    n = L.gbref_read_tracking_dump(str(path).encode(), out.reshape(-1), max_epochs)
    return n, out[: max(0, min(n, max_epochs))]


class RefAcqBlock:
    """The reference's pcps_acquisition block (compiled from /root/reference), driven dwell by dwell.
    Keyword arguments as OracleAcq / gb.Acquisition, plus system / signal / prn (GLONASS FDMA: signal '1G' and a PRN whose
    GLONASS_PRN channel number gives the offset) and use_fsm."""

    def __init__(self, system=b"G", signal=b"1C", prn=1, use_fsm=0, **kw):
        self.L = ref_blocks()
        d = dict(sampled_ms=1, ms_per_code=1, samples_per_chip=4, max_dwells=1, doppler_max=5000, doppler_step=250, num_doppler_bins_step2=4,
                 doppler_step2=125.0, fs_in=4000000, samples_per_ms=4000.0, samples_per_code=4000.0, bit_transition_flag=0,
                 use_CFAR_algorithm_flag=1, make_2_steps=0, cshort=0, threshold=0.0)
        d.update(kw)
        self.cshort = d["cshort"]
        self.conf = GbRefAcqConf(system=system, signal=signal, prn=prn, use_fsm=use_fsm, **d)
        self.h = self.L.gbref_acq_block_create(C.byref(self.conf))
        st = self.state()
        self.fft_size, self.consumed, self.num_bins = st.fft_size, st.consumed_samples, st.num_doppler_bins

    def state(self):
        s = GbRefAcqState()
        self.L.gbref_acq_block_get_state(self.h, C.byref(s))
        return s

    def set_local_code(self, code_c64):
        need = self.fft_size // 2 if self.conf.bit_transition_flag else self.consumed
        code = np.ascontiguousarray(code_c64, np.complex64)
        assert len(code) >= need
        self.L.gbref_acq_block_set_local_code(self.h, code.view(np.float32).copy())

    def reset(self):
        """AcquisitionInterface::reset() of the adapters = set_active(true)"""
        self.L.gbref_acq_block_set_active(self.h, 1)

    def start(self, sample_counter=0):
        """what the unit tests / ChannelFsm::start_acquisition do before the first dwell: set_state(1) (reset() = set_active(true))"""
        self.L.gbref_acq_block_set_sample_counter(self.h, sample_counter)
        self.L.gbref_acq_block_set_state(self.h, 1)

    def dwell(self, samples):
        """Feeds general_work() from `samples` (complex64, or interleaved int16 when cshort) until acquisition_core has run once;
        returns (GbRefAcqState, event) with event 1 positive / 2 negative / 0 none (more dwells or step two armed)."""
        buf = np.ascontiguousarray(samples, np.int16 if self.cshort else np.complex64)
        isz = 4 if self.cshort else 8
        n_items = buf.nbytes // isz
        pos, ev = 0, C.c_int(0)
        before = self.state()
        for _ in range(64):
            st0 = self.state()
            if not st0.active and not st0.step_two:
                raise RuntimeError("block is in standby: call start() (set_state(1)) or reset() first")
            # (standby with step two armed: general_work re-centres the fine grid and re-activates the block by itself, :978-985)
            was_state = st0.state
            used = self.L.gbref_acq_block_work(self.h, C.c_void_p(buf.ctypes.data + pos * isz), n_items - pos, C.byref(ev))
            pos += used
            if was_state == 2:
                return self.state(), ev.value
        raise RuntimeError("acquisition_core did not run (state %d -> %d)" % (before.state, self.state().state))

    def grid(self):
        eff = self.fft_size // 2 if self.conf.bit_transition_flag else self.fft_size
        return np.stack([np.ctypeslib.as_array(self.L.gbref_acq_block_grid_row(self.h, b), shape=(self.fft_size,))[:eff].copy() for b in range(self.num_bins)])

    def close(self):
        if self.h:
            self.L.gbref_acq_block_destroy(self.h)
            self.h = None


class RefTrkBlock:
    """The reference's dll_pll_veml_tracking block (compiled from /root/reference)."""

    def __init__(self, **kw):
        self.L = ref_blocks()
        d = dict(fs_in=4e6, vector_length=4000, system=b"G", signal=b"1C", pll_filter_order=3, dll_filter_order=2, fll_filter_order=1,
                 enable_fll_pull_in=0, enable_fll_steady_state=0, pull_in_time_s=2, fll_bw_hz=35.0, pll_bw_hz=50.0, dll_bw_hz=2.0, pll_bw_narrow_hz=20.0,
                 dll_bw_narrow_hz=2.0, early_late_space_chips=0.5, very_early_late_space_chips=0.5, early_late_space_narrow_chips=0.5,
                 very_early_late_space_narrow_chips=0.5, extend_correlation_symbols=1, high_dyn=0, smoother_length=10, cn0_samples=20, cn0_min=25,
                 max_lock_fail=50, carrier_lock_th=0.85, track_pilot=0, dump=0, dump_filename=b"")
        d.update(kw)
        self.conf = GbRefTrkConf(**d)
        self.h = self.L.gbref_trk_block_create(C.byref(self.conf))

    def start(self, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, first_item_index=0):
        self.L.gbref_trk_block_start(self.h, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, first_item_index)

    def run(self, samples_c64, max_epochs):
        x = np.ascontiguousarray(samples_c64, np.complex64).view(np.float32)
        out = (GbRefTrkEpoch * max_epochs)()
        ev = np.zeros(64, np.int32)
        nev = C.c_int(0)
        n = self.L.gbref_trk_block_run(self.h, x, len(x) // 2, out, max_epochs, ev, 64, C.byref(nev))
        return [out[i] for i in range(n)], list(ev[: nev.value])

    def state(self):
        return self.L.gbref_trk_block_state(self.h)

    def telemetry_fault(self):
        self.L.gbref_trk_block_telemetry_fault(self.h)

    def close(self):
        if self.h:
            self.L.gbref_trk_block_destroy(self.h)
            self.h = None


# ---------------------------------------------------------------- the reference's sibling blocks (oracle/ref_blocks2.cc)
class GbRefAcqvCall(C.Structure):
    _fields_ = [("state_before", C.c_int32), ("state_after", C.c_int32), ("consumed", C.c_int32), ("event", C.c_int32), ("test_statistics", C.c_float),
                ("input_power", C.c_float), ("mag", C.c_float), ("count", C.c_uint32), ("tong_count", C.c_uint32), ("acq_delay_samples", C.c_double),
                ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64), ("acq_doppler_step", C.c_uint32),
                ("active", C.c_int32)]


REF_TONG, REF_8MS, REF_E5A, REF_QUICKSYNC, REF_FINE_DOPPLER = 0, 1, 2, 3, 4


def ref_blocks2():
    L = ref_blocks()
    if L is None or not hasattr(L, "gbref_acqv_create"):
        return None
    if getattr(L, "_blocks2_bound", False):
        return L
    L.gbref_acqv_create.argtypes = [C.c_int, f64p, C.c_char, C.c_char_p]
    L.gbref_acqv_create.restype = C.c_void_p
    L.gbref_acqv_destroy.argtypes = [C.c_void_p]
    L.gbref_acqv_vector_items.argtypes = [C.c_void_p]
    L.gbref_acqv_vector_items.restype = C.c_uint32
    L.gbref_acqv_init.argtypes = [C.c_void_p]
    L.gbref_acqv_set_local_code.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.gbref_acqv_set_state.argtypes = [C.c_void_p, C.c_int]
    L.gbref_acqv_set_active.argtypes = [C.c_void_p, C.c_int]
    L.gbref_acqv_work.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(GbRefAcqvCall)]
    L.gbref_acqv_caf_vector.argtypes = [C.c_void_p, f32p, C.c_int]
    L.gbref_set_flags.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
    L._blocks2_bound = True
    return L


class RefAcqVariantBlock:
    """pcps_tong_acquisition_cc / galileo_pcps_8ms_acquisition_cc / galileo_e5a_noncoherentIQ_acquisition_caf_cc /
    pcps_quicksync_acquisition_cc compiled from /root/reference, one general_work() call at a time."""

    def __init__(self, variant, sampled_ms, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, max_dwells=1, tong_init_val=1,
                 tong_max_val=2, tong_max_dwells=20, bit_transition_flag=0, both_signal_components=0, caf_window_hz=0, zero_padding=0, folding_factor=2, prn=1,
                 system=b"G", signal=b"1C"):
        self.L = ref_blocks2()
        cfg = np.array([sampled_ms, max_dwells, doppler_max, doppler_step, fs_in, samples_per_ms, samples_per_code, threshold, tong_init_val, tong_max_val,
                        tong_max_dwells, bit_transition_flag, both_signal_components, caf_window_hz, zero_padding, folding_factor, prn], np.float64)
        self.h = self.L.gbref_acqv_create(variant, cfg, system, signal)
        self.vector_items = self.L.gbref_acqv_vector_items(self.h)
        self.L.gbref_acqv_init(self.h)
        self._keep = []

    def set_local_code(self, code_c64, code_q_c64=None):
        c = np.ascontiguousarray(code_c64, np.complex64).copy()
        q = np.ascontiguousarray(code_q_c64, np.complex64).copy() if code_q_c64 is not None else None
        self._keep = [c, q]
        self.L.gbref_acqv_set_local_code(self.h, c.ctypes.data_as(C.c_void_p), q.ctypes.data_as(C.c_void_p) if q is not None else None)

    def set_state(self, state):
        self.L.gbref_acqv_set_state(self.h, state)

    def set_active(self, active=True):
        self.L.gbref_acqv_set_active(self.h, int(active))

    def work(self, items_c64, n_in):
        x = np.ascontiguousarray(items_c64, np.complex64)
        r = GbRefAcqvCall()
        self.L.gbref_acqv_work(self.h, x.ctypes.data_as(C.c_void_p), n_in, C.byref(r))
        return r

    def caf_vector(self, n):
        out = np.zeros(n, np.float32)
        k = self.L.gbref_acqv_caf_vector(self.h, out, n)
        return out[:k]

    def close(self):
        if self.h:
            self.L.gbref_acqv_destroy(self.h)
            self.h = None


TRK2_FIELDS = ("tracking_sample_counter", "prompt_i", "prompt_q", "carrier_doppler_hz", "code_phase_samples", "carrier_phase_rads", "cn0_db_hz",
               "abs_e", "abs_p", "abs_l", "flags", "consumed", "correlation_length_ms")
REF_TRK_GLONASS, REF_TRK_CAID_CC, REF_TRK_CAID_SC, REF_TRK_KF = 0, 1, 2, 3


class RefSiblingTracker:
    """glonass_l1_ca_dll_pll_tracking_cc / gps_l1_ca_dll_pll_c_aid_tracking_cc / _sc compiled from /root/reference."""

    def __init__(self, kind, fs_in, vector_length, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz=20.0, dll_bw_narrow_hz=2.0, early_late_space_chips=0.5,
                 extend_correlation_ms=1, order=2, bce_run=False, bce_ptrans=0, bce_strans=0, bce_nu=0, bce_kappa=0):
        self.L = ref_blocks2()
        self.L.gbref_trk2_create.argtypes = [C.c_int, f64p]
        self.L.gbref_trk2_create.restype = C.c_void_p
        self.L.gbref_trk2_destroy.argtypes = [C.c_void_p]
        self.L.gbref_trk2_start.argtypes = [C.c_void_p, C.c_uint32, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        self.L.gbref_trk2_preamble.argtypes = [C.c_void_p, C.c_double]
        self.L.gbref_trk2_run.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, f64p]
        self.kind = kind
        self.L.gbref_trk2_set_doppler_step.argtypes = [C.c_void_p, C.c_uint32]
        conf = np.array([fs_in, vector_length, pll_bw_hz, dll_bw_hz, pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_chips, extend_correlation_ms,
                         order, 1.0 if bce_run else 0.0, bce_ptrans, bce_strans, bce_nu, bce_kappa], np.float64)
        self.h = self.L.gbref_trk2_create(kind, conf)

    def start(self, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, first_item_index=0, acq_doppler_step=0):
        self.L.gbref_trk2_set_doppler_step(self.h, acq_doppler_step)
        self.L.gbref_trk2_start(self.h, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, first_item_index)

    def preamble(self, timestamp_s):
        self.L.gbref_trk2_preamble(self.h, timestamp_s)

    def run(self, samples, n_calls):
        x = np.ascontiguousarray(samples, np.int16 if self.kind == REF_TRK_CAID_SC else np.complex64)
        n_items = x.size // 2 if self.kind == REF_TRK_CAID_SC else x.size
        rows = np.zeros((n_calls, 13), np.float64)
        n = self.L.gbref_trk2_run(self.h, x.ctypes.data_as(C.c_void_p), n_items, n_calls, rows.reshape(-1))
        return rows[:n]

    def close(self):
        if self.h:
            self.L.gbref_trk2_destroy(self.h)
            self.h = None
