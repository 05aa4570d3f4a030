"""ctypes binding of libgnss_b200.so (include/gnss_b200.h) -- thin, no compute in Python.

The library is the product; this module only marshals arguments.  It fails loudly when the shared
library is missing (`build()` in __graft_entry__.py makes it) and every compute call fails with
GB200_E_NO_DEVICE when there is no B200: there is no CPU fallback anywhere on this path.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgnss_b200.so")

ITEM_32FC, ITEM_16IC, ITEM_8IC = 0, 1, 2
CODE_REAL, CODE_CPLX, CODE_16SC = 0, 1, 2
KCLASS = {"trk": 0, "acq_fwd_col": 1, "acq_fwd_row": 2, "acq_inv_row": 3, "acq_inv_col": 4}
MAX_TAPS = 8
OUT_STRIDE = MAX_TAPS + 1
ITEM_DTYPES = {ITEM_32FC: np.complex64, ITEM_16IC: np.int16, ITEM_8IC: np.int8}

E_NAMES = {0: "OK", -1: "E_INVALID", -2: "E_NO_DEVICE", -3: "E_CUDA", -4: "E_RANGE", -5: "E_STATE", -6: "E_UNSUPPORTED"}


class Gb200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s: %s" % (E_NAMES.get(code, code), msg))
        self.code = code


class ServiceCfg(C.Structure):
    _fields_ = [("device", C.c_int), ("max_streams", C.c_int), ("ring_capacity", C.c_uint64), ("ring_slack", C.c_uint64),
                ("batch_window_us", C.c_uint32), ("reserved_", C.c_uint32)]


class AcqCfg(C.Structure):
    _fields_ = [
        ("sampled_ms", C.c_uint32), ("ms_per_code", C.c_uint32), ("samples_per_chip", C.c_uint32),
        ("max_dwells", C.c_uint32), ("doppler_max", C.c_uint32), ("doppler_step", C.c_uint32),
        ("num_doppler_bins_step2", C.c_uint32), ("doppler_step2", C.c_float), ("fs_in", C.c_int64),
        ("samples_per_ms", C.c_float), ("samples_per_code", C.c_float), ("bit_transition_flag", C.c_int),
        ("use_CFAR_algorithm_flag", C.c_int), ("make_2_steps", C.c_int), ("item_type", C.c_int),
        ("threshold", C.c_float), ("freq_offset_hz", C.c_int64),
    ]


class AcqResult(C.Structure):
    _fields_ = [
        ("code_phase_index", C.c_uint32), ("doppler_hz", C.c_int32), ("doppler_bin", C.c_uint32),
        ("peak_mag", C.c_float), ("input_power", C.c_float), ("test_statistic", C.c_float),
        ("decision", C.c_int), ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double),
        ("acq_samplestamp", C.c_uint64),
    ]


class AcqIqBin(C.Structure):
    """gb200_acq_iq_bin: one Doppler bin of gb200_acq_iq_combine."""
    _fields_ = [("peak", C.c_float * 4), ("index", C.c_uint32 * 4), ("ib_at_qb", C.c_float), ("mag", C.c_float),
                ("indext", C.c_uint32), ("picked", C.c_uint32)]


class TrkCfg(C.Structure):
    _fields_ = [("n_taps", C.c_int), ("max_code_len", C.c_int), ("code_kind", C.c_int), ("high_dyn", C.c_int), ("item_type", C.c_int)]


class TrkReq(C.Structure):
    _fields_ = [
        ("trk", C.c_void_p), ("stream_id", C.c_int), ("first_sample_index", C.c_uint64),
        ("rem_carr_phase_rad", C.c_float), ("phase_step_rad", C.c_float), ("phase_rate_step_rad", C.c_float),
        ("rem_code_phase_chips", C.c_float), ("code_phase_step_chips", C.c_float), ("code_phase_rate_step_chips", C.c_float),
        ("n_samples", C.c_int),
    ]


EXPORTS = [
    "gb200_service_create", "gb200_service_destroy", "gb200_last_error", "gb200_version", "gb200_service_cuda_stream",
    "gb200_service_synchronize", "gb200_service_launch_count", "gb200_service_timer_start", "gb200_service_timer_stop",
    "gb200_service_flush_l2", "gb200_service_coalescing_stats", "gb200_service_acq_coalescing_stats", "gb200_service_kernel_timing", "gb200_service_kernel_stats",
    "gb200_service_measure_fp32", "gb200_service_set_tuning", "gb200_stream_wait", "gb200_stream_push_shared", "gb200_stream_push", "gb200_stream_push_device", "gb200_stream_set_inverted_spectrum", "gb200_stream_set_decimator", "gb200_firdes_low_pass", "gb200_stream_read",
    "gb200_acq_create", "gb200_acq_create_folded", "gb200_acq_destroy", "gb200_acq_fft_size", "gb200_acq_consumed_samples", "gb200_acq_num_doppler_bins",
    "gb200_acq_set_local_code", "gb200_acq_set_doppler", "gb200_acq_set_doppler_grid", "gb200_acq_set_scaled_accumulation", "gb200_acq_set_threshold", "gb200_acq_reset", "gb200_acq_run",
    "gb200_acq_run_batch", "gb200_acq_launch_batch", "gb200_acq_collect_batch", "gb200_acq_discard_batch", "gb200_acq_keep_grid", "gb200_acq_read_grid", "gb200_acq_iq_combine", "gb200_acq_second_peak", "gb200_acq_wiped_spectrum_peak", "gb200_second_peak_noalloc",
    "gb200_trk_create", "gb200_trk_destroy", "gb200_trk_set_local_code_and_taps", "gb200_trk_set_data_code", "gb200_trk_set_taps", "gb200_trk_correlate",
    "gb200_trk_correlate_batch", "gb200_trk_correlate_16sc", "gb200_trk_correlate_16sc_batch", "gb200_trk_batch_create", "gb200_trk_batch_launch", "gb200_trk_batch_fetch", "gb200_trk_batch_destroy",
]

_lib = None


def lib():
    """Load libgnss_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libgnss_b200.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "or `make -C gnss-sdr-1_b200` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i, u32, u64, f = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_float
    L.gb200_service_create.argtypes = [C.POINTER(ServiceCfg), C.POINTER(vp)]
    L.gb200_service_destroy.argtypes = [vp]
    L.gb200_last_error.argtypes = [vp]
    L.gb200_last_error.restype = C.c_char_p
    L.gb200_version.restype = C.c_char_p
    L.gb200_service_cuda_stream.argtypes = [vp]
    L.gb200_service_cuda_stream.restype = vp
    L.gb200_service_synchronize.argtypes = [vp]
    L.gb200_service_launch_count.argtypes = [vp]
    L.gb200_service_launch_count.restype = u64
    L.gb200_service_timer_start.argtypes = [vp]
    L.gb200_service_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
    L.gb200_service_flush_l2.argtypes = [vp, C.c_size_t]
    L.gb200_service_coalescing_stats.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.gb200_service_acq_coalescing_stats.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.gb200_service_measure_fp32.argtypes = [vp, C.POINTER(C.c_double)]
    L.gb200_service_kernel_timing.argtypes = [vp, i]
    L.gb200_service_kernel_stats.argtypes = [vp, i, C.POINTER(C.c_double), C.POINTER(u64)]
    L.gb200_stream_push.argtypes = [vp, i, vp, C.c_size_t, i, u64]
    L.gb200_service_set_tuning.argtypes = [vp, C.c_char_p, i]
    L.gb200_stream_wait.argtypes = [vp, i]
    L.gb200_stream_push_shared.argtypes = [vp, i, vp, C.c_size_t, i, u64]
    L.gb200_stream_set_decimator.argtypes = [vp, i, i, u32, vp, u32]
    L.gb200_firdes_low_pass.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, vp, C.POINTER(u32)]
    L.gb200_stream_read.argtypes = [vp, i, u64, C.c_size_t, vp]
    L.gb200_stream_push_device.argtypes = [vp, i, vp, C.c_size_t, i, u64]
    L.gb200_acq_create.argtypes = [vp, C.POINTER(AcqCfg), C.POINTER(vp)]
    L.gb200_acq_create_folded.argtypes = [vp, C.POINTER(AcqCfg), C.c_uint32, C.POINTER(vp)]
    L.gb200_acq_destroy.argtypes = [vp]
    for fn in (L.gb200_acq_fft_size, L.gb200_acq_consumed_samples, L.gb200_acq_num_doppler_bins):
        fn.argtypes = [vp]
        fn.restype = u32
    L.gb200_acq_set_local_code.argtypes = [vp, vp, C.c_size_t]
    L.gb200_acq_set_doppler.argtypes = [vp, u32, u32, C.c_int64]
    L.gb200_acq_set_threshold.argtypes = [vp, f]
    L.gb200_acq_reset.argtypes = [vp]
    L.gb200_acq_set_doppler_grid.argtypes = [vp, u32, vp]
    L.gb200_acq_set_scaled_accumulation.argtypes = [vp, i]
    L.gb200_acq_run.argtypes = [vp, i, u64, C.POINTER(AcqResult)]
    L.gb200_acq_run_batch.argtypes = [C.POINTER(vp), i, i, u64, C.POINTER(AcqResult)]
    L.gb200_acq_launch_batch.argtypes = [C.POINTER(vp), i, i, u64]
    L.gb200_acq_collect_batch.argtypes = [C.POINTER(vp), i, C.POINTER(AcqResult)]
    L.gb200_acq_discard_batch.argtypes = [C.POINTER(vp), i]
    L.gb200_acq_keep_grid.argtypes = [vp, i]
    L.gb200_acq_read_grid.argtypes = [vp, vp]
    L.gb200_acq_iq_combine.argtypes = [vp, vp, vp, vp, f, C.POINTER(AcqIqBin)]
    L.gb200_acq_second_peak.argtypes = [vp, u32, u32, u32, C.POINTER(C.c_float)]
    L.gb200_second_peak_noalloc.argtypes = [vp, u32, u32, u32, vp, vp, C.c_char_p, C.c_size_t]
    L.gb200_acq_wiped_spectrum_peak.argtypes = [vp, i, u64, vp, u32, u32, C.POINTER(u32), C.POINTER(C.c_float)]
    L.gb200_trk_create.argtypes = [vp, C.POINTER(TrkCfg), C.POINTER(vp)]
    L.gb200_trk_destroy.argtypes = [vp]
    L.gb200_trk_set_local_code_and_taps.argtypes = [vp, vp, i, vp, i]
    L.gb200_trk_set_data_code.argtypes = [vp, vp, i, i]
    L.gb200_trk_correlate.argtypes = [vp, i, u64, f, f, f, f, f, f, i, vp]
    L.gb200_trk_correlate_batch.argtypes = [vp, i, C.POINTER(TrkReq), vp]
    L.gb200_trk_correlate_16sc.argtypes = [vp, i, u64, f, f, f, f, i, vp]
    L.gb200_trk_correlate_16sc_batch.argtypes = [vp, i, C.POINTER(TrkReq), vp]
    L.gb200_trk_batch_create.argtypes = [vp, i, C.POINTER(TrkReq), C.POINTER(vp)]
    L.gb200_trk_batch_launch.argtypes = [vp]
    L.gb200_trk_batch_fetch.argtypes = [vp, vp]
    L.gb200_trk_batch_destroy.argtypes = [vp]
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Service:
    """gb200_service: one per process and GPU."""

    def __init__(self, device=-1, max_streams=4, ring_capacity=1 << 24, ring_slack=1 << 21, batch_window_us=0):
        self.L = lib()
        self.h = C.c_void_p()
        cfg = ServiceCfg(device, max_streams, ring_capacity, ring_slack, batch_window_us, 0)
        rc = self.L.gb200_service_create(C.byref(cfg), C.byref(self.h))
        if rc:
            raise Gb200Error(rc, self.L.gb200_last_error(None).decode())

    def check(self, rc):
        if rc:
            raise Gb200Error(rc, self.L.gb200_last_error(self.h).decode())

    def set_tuning(self, key, value):
        """profiling / experiment switches (gb200_service_set_tuning)"""
        self.check(self.L.gb200_service_set_tuning(self.h, key.encode(), int(value)))

    def tune_from_env(self, environ=None):
        """GB200_TUNE="acq_lanes=1,acq_fused=1": the profiling tools' way of steering a bench run (never read by the library)"""
        spec = (environ if environ is not None else os.environ).get("GB200_TUNE", "")
        for kv in filter(None, spec.split(",")):
            k, v = kv.split("=")
            self.set_tuning(k.strip(), int(v))
        return spec

    def push_shared(self, stream_id, samples, item_type, first_sample_index):
        """one of several producers of the same stream: only items beyond the ring's newest are copied"""
        a = np.ascontiguousarray(samples, ITEM_DTYPES[item_type])
        n_items = a.size if item_type == ITEM_32FC else a.size // 2
        self.check(self.L.gb200_stream_push_shared(self.h, stream_id, _ptr(a), n_items, item_type, first_sample_index))

    def wait_pushes(self, stream_id):
        self.check(self.L.gb200_stream_wait(self.h, stream_id))

    def coalescing_stats(self):
        b, r = C.c_uint64(), C.c_uint64()
        self.check(self.L.gb200_service_coalescing_stats(self.h, C.byref(b), C.byref(r)))
        return b.value, r.value

    def acq_coalescing_stats(self):
        b, r = C.c_uint64(), C.c_uint64()
        self.check(self.L.gb200_service_acq_coalescing_stats(self.h, C.byref(b), C.byref(r)))
        return b.value, r.value

    def close(self):
        if self.h:
            self.L.gb200_service_destroy(self.h)
            self.h = C.c_void_p()

    @property
    def cuda_stream(self):
        return self.L.gb200_service_cuda_stream(self.h)

    def synchronize(self):
        self.check(self.L.gb200_service_synchronize(self.h))

    def launch_count(self):
        return int(self.L.gb200_service_launch_count(self.h))

    def timer_start(self):
        self.check(self.L.gb200_service_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self.check(self.L.gb200_service_timer_stop(self.h, C.byref(ms)))
        return float(ms.value)

    def measure_fp32_tflops(self):
        v = C.c_double()
        self.check(self.L.gb200_service_measure_fp32(self.h, C.byref(v)))
        return float(v.value)

    def kernel_timing(self, enable=True):
        self.check(self.L.gb200_service_kernel_timing(self.h, int(enable)))

    def kernel_stats(self, kernel_class):
        ms, n = C.c_double(), C.c_uint64()
        self.check(self.L.gb200_service_kernel_stats(self.h, kernel_class, C.byref(ms), C.byref(n)))
        return float(ms.value), int(n.value)

    def flush_l2(self, nbytes=256 << 20):
        self.check(self.L.gb200_service_flush_l2(self.h, nbytes))

    def push(self, stream_id, samples, item_type, first_sample_index):
        """samples: numpy complex64 (32FC) or interleaved int16 / int8 (I,Q,I,Q...)."""
        a = np.ascontiguousarray(samples, ITEM_DTYPES[item_type])
        n_items = a.size if item_type == ITEM_32FC else a.size // 2
        self.check(self.L.gb200_stream_push(self.h, stream_id, _ptr(a), n_items, item_type, first_sample_index))

    def set_inverted_spectrum(self, stream_id, inverted):
        """inverted_spectrum of the data-type adapters: later pushes to this stream are conjugated on the device."""
        self.check(self.L.gb200_stream_set_inverted_spectrum(self.h, stream_id, int(inverted)))

    def set_decimator(self, src_stream, dst_stream, decimation, taps=None):
        """acquisition resampler (gnss_flowgraph.cc:423-476): later pushes to src_stream also feed the FIR-decimated items to dst_stream"""
        t = np.ascontiguousarray(taps if taps is not None else [1.0], np.float32)
        self.check(self.L.gb200_stream_set_decimator(self.h, src_stream, dst_stream, decimation, _ptr(t), len(t)))

    def read(self, stream_id, first_sample_index, n_items, item_type=ITEM_32FC):
        """items of a stream back on the host, in their wire type"""
        out = np.zeros(n_items if item_type == ITEM_32FC else 2 * n_items, ITEM_DTYPES[item_type])
        self.check(self.L.gb200_stream_read(self.h, stream_id, first_sample_index, n_items, _ptr(out)))
        return out

    def push_ptr(self, stream_id, host_ptr, n_items, item_type, first_sample_index):
        self.check(self.L.gb200_stream_push(self.h, stream_id, C.c_void_p(host_ptr), n_items, item_type, first_sample_index))

    def push_device_ptr(self, stream_id, dev_ptr, n_items, item_type, first_sample_index):
        self.check(self.L.gb200_stream_push_device(self.h, stream_id, C.c_void_p(dev_ptr), n_items, item_type, first_sample_index))


class Acquisition:
    """gb200_acq: one pcps_acquisition block instance (one channel / PRN)."""

    def __init__(self, service, fold=1, **kw):
        self.s = service
        d = dict(sampled_ms=1, ms_per_code=1, samples_per_chip=4, max_dwells=1, doppler_max=5000, doppler_step=250,
                 num_doppler_bins_step2=4, doppler_step2=125.0, fs_in=4000000, samples_per_ms=4000.0, samples_per_code=4000.0,
                 bit_transition_flag=0, use_CFAR_algorithm_flag=1, make_2_steps=0, item_type=ITEM_32FC, threshold=0.0,
                 freq_offset_hz=0)
        d.update(kw)
        self.cfg = AcqCfg(**d)
        self.h = C.c_void_p()
        # fold > 1: QuickSync geometry (the wiped block folded `fold` times onto consumed / fold FFT points)
        service.check(service.L.gb200_acq_create_folded(service.h, C.byref(self.cfg), fold, C.byref(self.h)))
        L = service.L
        self.fft_size = L.gb200_acq_fft_size(self.h)
        self.consumed = L.gb200_acq_consumed_samples(self.h)
        self.num_bins = L.gb200_acq_num_doppler_bins(self.h)

    def set_local_code(self, code_c64):
        a = np.ascontiguousarray(code_c64, np.complex64)
        self.s.check(self.s.L.gb200_acq_set_local_code(self.h, _ptr(a), a.size))

    def set_doppler(self, doppler_max, doppler_step, freq_offset_hz=0):
        self.s.check(self.s.L.gb200_acq_set_doppler(self.h, doppler_max, doppler_step, freq_offset_hz))
        self.num_bins = self.s.L.gb200_acq_num_doppler_bins(self.h)

    def set_doppler_grid(self, phase_steps_rad):
        """explicit per-bin phase steps (sibling PCPS blocks with their own grid rule)"""
        st = np.ascontiguousarray(phase_steps_rad, np.float32)
        self.s.check(self.s.L.gb200_acq_set_doppler_grid(self.h, len(st), _ptr(st)))
        self.num_bins = self.s.L.gb200_acq_num_doppler_bins(self.h)

    def set_scaled_accumulation(self, enable=True):
        """Tong detector arithmetic: grid += |y|^2 / (N^4 input_power) every dwell"""
        self.s.check(self.s.L.gb200_acq_set_scaled_accumulation(self.h, 1 if enable else 0))
        self.num_bins = self.s.L.gb200_acq_num_doppler_bins(self.h)

    def set_threshold(self, th):
        self.s.check(self.s.L.gb200_acq_set_threshold(self.h, th))

    def reset(self):
        self.s.check(self.s.L.gb200_acq_reset(self.h))

    def keep_grid(self, keep=True):
        self.s.check(self.s.L.gb200_acq_keep_grid(self.h, int(keep)))

    def run(self, stream_id, first_sample_index):
        r = AcqResult()
        self.s.check(self.s.L.gb200_acq_run(self.h, stream_id, first_sample_index, C.byref(r)))
        return r

    def read_grid(self):
        eff = self.fft_size // 2 if self.cfg.bit_transition_flag else self.fft_size
        g = np.zeros((self.num_bins, eff), np.float32)
        self.s.check(self.s.L.gb200_acq_read_grid(self.h, _ptr(g)))
        return g

    def close(self):
        if self.h:
            self.s.L.gb200_acq_destroy(self.h)
            self.h = C.c_void_p()


def firdes_low_pass(gain, sampling_freq, cutoff_freq, transition_width):
    """gr::filter::firdes::low_pass(.., WIN_HAMMING) as the C ABI restates it (host arithmetic)"""
    L = lib()
    n = C.c_uint32(0)
    one = np.zeros(1, np.float32)
    rc = L.gb200_firdes_low_pass(gain, sampling_freq, cutoff_freq, transition_width, _ptr(one), C.byref(n))
    if rc != -4:
        raise Gb200Error(rc, "gb200_firdes_low_pass")
    taps = np.zeros(n.value, np.float32)
    rc = L.gb200_firdes_low_pass(gain, sampling_freq, cutoff_freq, transition_width, _ptr(taps), C.byref(n))
    if rc:
        raise Gb200Error(rc, "gb200_firdes_low_pass")
    return taps


def acq_iq_combine(ia, qa=None, ib=None, qb=None, norm=1.0):
    """gb200_acq_iq_combine over the grids the handles kept from their last batch; returns the per-bin records."""
    rec = (AcqIqBin * ia.num_bins)()
    h = [a.h if a is not None else None for a in (ia, qa, ib, qb)]
    ia.s.check(ia.s.L.gb200_acq_iq_combine(h[0], h[1], h[2], h[3], norm, rec))
    return rec


def acq_handles(acqs):
    return (C.c_void_p * len(acqs))(*[a.h for a in acqs])


def acq_run_batch(acqs, stream_id, first_sample_index):
    s = acqs[0].s
    res = (AcqResult * len(acqs))()
    s.check(s.L.gb200_acq_run_batch(acq_handles(acqs), len(acqs), stream_id, first_sample_index, res))
    return list(res)


def acq_launch_batch(acqs, handles, stream_id, first_sample_index):
    s = acqs[0].s
    s.check(s.L.gb200_acq_launch_batch(handles, len(acqs), stream_id, first_sample_index))


def acq_discard_batch(acqs, handles):
    """stop_acquisition() for a launched batch: abandon it without waiting"""
    s = acqs[0].s
    s.check(s.L.gb200_acq_discard_batch(handles, len(acqs)))


def acq_collect_batch(acqs, handles):
    s = acqs[0].s
    res = (AcqResult * len(acqs))()
    s.check(s.L.gb200_acq_collect_batch(handles, len(acqs), res))
    return list(res)


class Multicorrelator:
    """gb200_trk: one Cpu_Multicorrelator[_Real_Codes] instance (one tracking channel)."""

    def __init__(self, service, n_taps, max_code_len, code_kind=CODE_REAL, high_dyn=False, item_type=ITEM_32FC):
        self.s = service
        self.cfg = TrkCfg(n_taps, max_code_len, code_kind, int(high_dyn), item_type)
        self.h = C.c_void_p()
        service.check(service.L.gb200_trk_create(service.h, C.byref(self.cfg), C.byref(self.h)))
        self.n_out = n_taps

    def set_local_code_and_taps(self, code, shifts):
        n_items = None
        if self.cfg.code_kind == CODE_CPLX:
            c = np.ascontiguousarray(code, np.complex64)
        elif self.cfg.code_kind == CODE_16SC:
            c = np.ascontiguousarray(code, np.int16)  # interleaved I, Q
            n_items = c.size // 2
        else:
            c = np.ascontiguousarray(code, np.float32)
        sh = np.ascontiguousarray(shifts, np.float32)
        self.s.check(self.s.L.gb200_trk_set_local_code_and_taps(self.h, _ptr(c), c.size if n_items is None else n_items, _ptr(sh), sh.size))
        self.n_taps = sh.size
        self.n_out = sh.size

    def set_taps(self, shifts):
        """Tap offsets only (narrow-spacing switch, dll_pll_veml_tracking.cc:1753-1764)."""
        sh = np.ascontiguousarray(shifts, np.float32)
        self.s.check(self.s.L.gb200_trk_set_taps(self.h, _ptr(sh), sh.size))

    def set_data_code(self, code, prompt_tap):
        c = np.ascontiguousarray(code, np.float32)
        self.s.check(self.s.L.gb200_trk_set_data_code(self.h, _ptr(c), c.size, prompt_tap))
        self.n_out = self.n_taps + 1

    def correlate(self, stream_id, first_sample_index, rem_carr, phase_step, rem_code, code_step, n_samples,
                  phase_rate=0.0, code_rate=0.0):
        out = np.zeros(self.n_out, np.complex64)
        self.s.check(self.s.L.gb200_trk_correlate(self.h, stream_id, first_sample_index, rem_carr, phase_step, phase_rate,
                                                  rem_code, code_step, code_rate, n_samples, _ptr(out)))
        return out

    def correlate_16sc(self, stream_id, first_sample_index, rem_carr, phase_step, rem_code, code_step, n_samples):
        """Cpu_Multicorrelator_16sc semantics; returns int16[n_taps, 2] (I, Q)."""
        out = np.zeros(2 * self.n_taps, np.int16)
        self.s.check(self.s.L.gb200_trk_correlate_16sc(self.h, stream_id, first_sample_index, rem_carr, phase_step, rem_code, code_step,
                                                       n_samples, _ptr(out)))
        return out.reshape(-1, 2)

    def close(self):
        if self.h:
            self.s.L.gb200_trk_destroy(self.h)
            self.h = C.c_void_p()


def correlate_16sc_batch(service, reqs):
    out = np.zeros((len(reqs), 2 * MAX_TAPS), np.int16)
    service.check(service.L.gb200_trk_correlate_16sc_batch(service.h, len(reqs), reqs, _ptr(out)))
    return out.reshape(len(reqs), MAX_TAPS, 2)


def make_reqs(items):
    """items: iterable of (Multicorrelator, stream_id, first, rem_carr, phase_step, phase_rate, rem_code, code_step, code_rate, n)."""
    items = list(items)
    arr = (TrkReq * len(items))()
    for k, (m, sid, first, rc, ps, pr, rcode, cs, cr, n) in enumerate(items):
        arr[k] = TrkReq(m.h, sid, first, rc, ps, pr, rcode, cs, cr, n)
    return arr


class TrkBatch:
    """gb200_trk_batch: staged requests that can be launched repeatedly."""

    def __init__(self, service, reqs):
        self.s = service
        self.n = len(reqs)
        self.h = C.c_void_p()
        service.check(service.L.gb200_trk_batch_create(service.h, self.n, reqs, C.byref(self.h)))

    def launch(self):
        self.s.check(self.s.L.gb200_trk_batch_launch(self.h))

    def fetch(self):
        out = np.zeros((self.n, OUT_STRIDE), np.complex64)
        self.s.check(self.s.L.gb200_trk_batch_fetch(self.h, _ptr(out)))
        return out

    def close(self):
        if self.h:
            self.s.L.gb200_trk_batch_destroy(self.h)
            self.h = C.c_void_p()


def correlate_batch(service, reqs):
    out = np.zeros((len(reqs), OUT_STRIDE), np.complex64)
    service.check(service.L.gb200_trk_correlate_batch(service.h, len(reqs), reqs, _ptr(out)))
    return out
