#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 GNSS correlator engine (contract: see the task brief).

Workload (BASELINE.json configs[1], "C2"): Galileo E1 B/C, 8 Msps interleaved-int16 input, 32 PRNs.
  step = ONE pass of the hot path over one batch:
    acquisition  32 PRN x 80 Doppler bins (+-5 kHz / 125 Hz) x 32 000 code phases (4 ms PCPS, CFAR)   = 81.92 M grid cells
    tracking     32 channels x 248 epochs x 32 000 samples, 5 taps (VE,E,P,L,VL) + data-prompt tap     = 253.95 M channel.samples
BASELINE.json's metric names both paths.  The top-level `value` is acquisition grid cells/s; tracking channel*Msamples/s is mirrored
at the top level as `tracking_value` / `tracking_e2e` (same keys in the reference arm) and detailed under `tracking`.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--mode weak|shard] [--sustain S] [--no-configs]

Numbers in one line:
  value / tracking_value   burst: CUDA events around each step, inputs resident in HBM, L2 flushed before every timed region
  e2e / tracking_e2e       the same through the C ABI with HOST buffers (H2D of the step's samples and D2H of its results inside)
  sustained                >= S seconds (default 2) of back-to-back steps per path, no host synchronisation or L2 flush inside,
                           clocks sampled throughout (in-process NVML): the number a long-running receiver sees
  config_results           C1 / C3 / C5 acquisition and tracking throughput on their own shapes (N = 1 only)
  parity                   the bench's own C2 batch checked against the CPU oracle (untimed)
--mode weak (default): every rank runs the C2 workload on its own PRNs / channels (no data-path collective).
--mode shard: ONE fixed search / channel set (C5 cold start: 32 GPS + 36 Galileo PRNs at 50 Msps; C4: 1024 channels) partitioned by
  gnss-sdr-1_b200/shard.py, with the NCCL gather of the peak records / correlator outputs and the tie-break merge INSIDE the
  timed region: "scaling": "strong" (bench_shard.py).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "gnss-sdr-1_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

FS = 8_000_000
N_FFT = 32_000          # 4 ms at 8 Msps
N_PRN = 32
DOPPLER_MAX, DOPPLER_STEP = 5000, 125
N_BINS = 80
ACQ_THRESHOLD = 7e-4     # CFAR statistic: noise-only maxima of the 2.56 M-cell grid reach ~6e-4, a 40 dB-Hz E1B reaches 9e-4
N_CH = 32
N_EPOCH = 248
EPOCH = 32_000
STREAM_ITEMS = 8_000_000  # 1 s
E2E_BLOCKS = 8             # the end-to-end tracking pass hands the second of samples over in this many blocks
TAPS = np.array([-1.2, -0.3, 0.0, 0.3, 1.2], np.float32)  # +-0.6 / +-0.15 chip on the 2-samples-per-chip E1 table
METRIC = "acq grid cells/s (tracking channel*Msamples/s in 'tracking_value' / 'tracking')"
WORKLOAD = ("C2: Galileo E1 4 ms PCPS acquisition (32 PRN x 80 bins x 32000 points, CFAR) + 5-tap VEML tracking with data-prompt tap "
            "(32 ch x 248 epochs x 32000 samples), 8 Msps interleaved int16, per GPU")


def traffic_constants():
    """DRAM bytes per launch of the dominant kernels, from the committed ncu summaries (profiles/r02_traffic.json is written by
    profiles/tools/traffic_from_ncu.py out of the `ncu --set full` capture of this bench; nothing is hard-coded here)."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(p):
        return json.load(open(p))
    return {"acq_pair_dram_bytes": None, "trk_dram_bytes": None, "source": "profiles/r02_traffic.json missing"}


def build_stream(seed, items=None):
    """Synthetic C2 stream: 10 E1 satellites (B - C), CN0 44..35 dB-Hz, int16 at 2^10 LSB per sigma.
    `items` < STREAM_ITEMS returns the identical prefix (noise is drawn in 1 M-item blocks)."""
    import synth

    sats = [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=1 + 3 * k, cn0_dbhz=44 - k, doppler_hz=-4500 + 1000 * k,
                 delay_samples=1000 + 2917 * k) for k in range(10)]
    rng = np.random.default_rng(seed)
    # noise in float32 blocks (8 M complex samples); signals added block-wise to bound memory
    total = STREAM_ITEMS if items is None else -(-int(items) // 1_000_000) * 1_000_000
    out = np.empty(2 * total, np.int16)
    blk = 1_000_000
    for b0 in range(0, total, blk):
        n = min(blk, total - b0)
        x = (rng.standard_normal(n, np.float32) + 1j * rng.standard_normal(n, np.float32)).astype(np.complex64)
        t = (b0 + np.arange(n, dtype=np.float64)) / FS
        for s in sats:
            tb = synth.code_table(s["kind"], s["prn"])
            tc = synth.code_table(s["pilot_kind"], s["prn"])
            rate = synth.TABLE_RATE[s["kind"]] * (1.0 + s["doppler_hz"] / synth.CARRIER_HZ[s["kind"]])
            idx = np.floor((t - s["delay_samples"] / FS) * rate).astype(np.int64) % len(tb)
            amp = np.sqrt(2.0 * 10.0 ** (s["cn0_dbhz"] / 10.0) / FS)
            x += (amp * (tb[idx] - tc[idx]) / np.sqrt(2.0) * np.exp(2j * np.pi * s["doppler_hz"] * t)).astype(np.complex64)
        out[2 * b0: 2 * (b0 + n)] = synth.quantize(x, "16ic")
    return out, sats


def acq_codes():
    import synth

    return [synth.sampled_code(synth.code_table("gal_e1b_sinboc", p), synth.TABLE_RATE["gal_e1b_sinboc"], FS, N_FFT) for p in range(1, N_PRN + 1)]


def trk_requests(rank, n_ch=N_CH, n_epoch=N_EPOCH, epoch=EPOCH, fs=FS, table_len=8184.0, carrier=1575.42e6):
    """Open-loop replay: per channel a Doppler, a start offset and slowly advancing remnant phases (float32 as the block passes them)."""
    reqs = []
    for c in range(n_ch):
        dop = -4500.0 + (290.0 if n_ch == N_CH else 9000.0 / max(1, n_ch)) * c + 7.0 * rank
        phase_step = np.float32(2 * np.pi * dop / fs)
        code_step = np.float32(table_len / epoch * (1.0 + dop / carrier))
        first0 = 1000 + 977 * (c % 32)
        for e in range(n_epoch):
            rem_carr = np.float32(np.fmod(float(phase_step) * epoch * e, 2 * np.pi))
            rem_code = np.float32(0.37 + 1e-3 * e + 0.01 * (c % 32))
            reqs.append((c, first0 + e * epoch, float(rem_carr), float(phase_step), float(rem_code), float(code_step)))
    return reqs


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons sampled DURING the timed regions (B200_PROFILING.md), in-process through NVML (nvidia_ml_py) so
    that sampling costs no process spawns; falls back to nvidia-smi.  One sampler, on rank 0 only."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.rows = []          # (t, sm_mhz, reasons bitmask, power_w)
        self.stop_flag = False
        self.nv = None
        self.h = None
        self.max_mhz = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag:
            try:
                if self.nv is not None:
                    nv = self.nv
                    mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    try:
                        reasons = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                    except Exception:
                        reasons = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                    try:
                        power = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                    except Exception:
                        power = None
                    self.rows.append((time.perf_counter(), mhz, reasons, power))
                else:
                    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
                    o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                       capture_output=True, text=True, timeout=5).stdout.strip()
                    c = [v.strip() for v in o.split(",")]
                    if len(c) >= 6 and c[0].replace(".", "").isdigit():
                        self.max_mhz = float(c[1])
                        bits = sum(b for b, v in zip((0x8, 0x40, 0x20, 0x4), c[2:6]) if v.lower().startswith("active"))
                        self.rows.append((time.perf_counter(), float(c[0]), bits, None))
            except Exception:
                pass
            time.sleep(self.period)

    def summary(self, t0=None, t1=None):
        rows = [r for r in self.rows if (t0 is None or r[0] >= t0) and (t1 is None or r[0] <= t1)]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no samples"], "samples": 0}
        sm = sorted(r[1] for r in rows)
        # NVML clocks-event-reason bits: 0x4 sw_power_cap, 0x8 hw_slowdown, 0x20 sw_thermal_slowdown, 0x40 hw_thermal_slowdown
        names = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))
        reasons = [n for b, n in names if any(r[2] & b for r in rows)]
        out = {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(rows),
               "source": "nvml" if self.nv is not None else "nvidia-smi"}
        pw = [r[3] for r in rows if r[3] is not None]
        if pw:
            out["power_w_max"] = max(pw)
        return out


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ============================================================================ CPU side (oracle / reference; bounded samples)
def cpu_tracking_baseline(seconds_target=8.0):
    """The reference's OWN code (oracle/_ref: Cpu_Multicorrelator_Real_Codes + volk_gnsssdr _u_avx protokernels compiled
    from /root/reference), one std::thread per channel on all host cores -- the shape of
    cpu_multicorrelator_real_codes_test.cc:133-163 with the block's default kernel pair.  Falls back to the C oracle
    port (scalar, 1 thread per core via processes is not attempted) only if oracle/_ref was never built."""
    import oracle_lib as ol
    import synth

    cores = os.cpu_count() or 1
    rng = np.random.default_rng(1)
    x = (rng.standard_normal(EPOCH) + 1j * rng.standard_normal(EPOCH)).astype(np.complex64).view(np.float32)
    code = synth.code_table("gal_e1c_sinboc", 1)
    dcode = synth.code_table("gal_e1b_sinboc", 1)
    R = ol.ref()
    if R is not None:
        iters = 8
        t = R.gbref_bench_multicorrelator_real(x, EPOCH, code, len(code), TAPS, len(TAPS), cores, iters, 1, dcode.ctypes.data)
        iters = max(8, int(iters * seconds_target / max(t, 1e-3)))
        t = R.gbref_bench_multicorrelator_real(x, EPOCH, code, len(code), TAPS, len(TAPS), cores, iters, 1, dcode.ctypes.data)
        v = cores * iters * EPOCH / t / 1e6
        t1 = R.gbref_bench_multicorrelator_real(x, EPOCH, code, len(code), TAPS, len(TAPS), 1, max(4, iters // 4), 1, dcode.ctypes.data)
        v1 = max(4, iters // 4) * EPOCH / t1 / 1e6
        return {"value": v, "unit": "channel*Msamples/s", "cores": cores, "kind": "reference", "value_1_core": v1,
                "sample": "%d threads x %d epochs x %d samples, 5 taps + data-prompt tap, reference volk_gnsssdr _u_avx resampler+rotator (%.1f s)" % (cores, iters, EPOCH, t)}
    O = ol.oracle()
    out = np.zeros(2 * len(TAPS), np.float32)
    n_it, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds_target:
        O.go_multicorrelator_real(x, EPOCH, code, len(code), TAPS, len(TAPS), 0.1, 0.0123, 0.0, 0.25, 8184.0 / EPOCH, 0.0, 0, out)
        n_it += 1
    t = time.perf_counter() - t0
    return {"value": n_it * EPOCH / t / 1e6, "unit": "channel*Msamples/s", "cores": 1, "kind": "port",
            "sample": "%d epochs x %d samples, scalar C restatement (oracle/_ref absent)" % (n_it, EPOCH)}


def _acq_worker(args):
    """One acquisition CHANNEL the way the reference runs it: a single-threaded pcps_acquisition per channel (the GNU Radio
    scheduler gives every block its own thread), here one worker process per channel on its own core."""
    x, code, seconds, fs, n_fft, dmax, dstep = args
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import acq_port

    port = acq_port.AcqPort(fs, n_fft, n_fft, dmax, dstep, workers=1)
    cf = [port.set_local_code(code)]
    res = port.run(x, cf)  # warm-up (plans, page faults)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        res = port.run(x, cf)
        n += 1
    return n, time.perf_counter() - t0, res[0]


def cpu_acquisition_baseline(stream_i16, codes, seconds_target=10.0):
    """Restated acquisition_core (oracle/acq_port.py: numpy element-wise passes + scipy.fft pocketfft float32 standing in for FFTW),
    CHANNEL-PARALLEL like the reference: `cores` single-threaded worker processes, one channel (PRN) each, all running for the
    same wall time.  "kind": "port" (FFTW-class CPU FFT, not FFTW).  Independent of OMP_NUM_THREADS / torchrun's thread caps."""
    import multiprocessing as mp

    cores = os.cpu_count() or 1
    x = (stream_i16[0: 2 * N_FFT: 2].astype(np.float32) + 1j * stream_i16[1: 2 * N_FFT: 2].astype(np.float32)).astype(np.complex64)
    ctx = mp.get_context("spawn")
    t0 = time.perf_counter()
    env_keep = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS")}
    for k in env_keep:
        os.environ[k] = "1"  # one thread per worker process, whatever the launcher set
    try:
        with ctx.Pool(cores) as pool:
            out = pool.map(_acq_worker, [(x, codes[k % len(codes)], seconds_target, FS, N_FFT, DOPPLER_MAX, DOPPLER_STEP) for k in range(cores)])
    finally:
        for k, v in env_keep.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    wall = time.perf_counter() - t0
    dwells = sum(o[0] for o in out)
    t = max(o[1] for o in out)
    cells = dwells * N_BINS * N_FFT
    return {"value": cells / t, "unit": "grid cells/s", "cores": cores, "kind": "port",
            "sample": "%d worker processes (one channel each, single-threaded) x %.1f s = %d PRN-dwells x %d bins x %d points, numpy + scipy.fft float32 "
                      "(FFTW-class CPU FFT, not FFTW); wall incl. process start %.1f s" % (cores, t, dwells, N_BINS, N_FFT, wall),
            "check": {"code_phase": out[0][2][0], "doppler_hz": out[0][2][1]}}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    stream, _ = build_stream(20260002, items=1_000_000)
    codes = acq_codes()
    steps = max(1, args.steps)
    acq_vals, trk_vals = [], []
    t0 = time.perf_counter()
    per = max(1.0, min(8.0, 50.0 / steps))
    for _ in range(steps):
        a = cpu_acquisition_baseline(stream, codes, per)
        t = cpu_tracking_baseline(per)
        acq_vals.append(a["value"])
        trk_vals.append(t["value"])
    wall = time.perf_counter() - t0
    a["value"] = float(np.mean(acq_vals))
    t["value"] = float(np.mean(trk_vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": a["value"], "unit": "grid cells/s", "n_gpus": world, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * wall / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "timing": "host wall clock, bounded sample per step"},
        "cpu_baseline": a,
        "e2e": {"value": a["value"], "unit": "grid cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "tracking_value": t["value"], "tracking_unit": "channel*Msamples/s", "tracking_e2e": t["value"],
        "tracking": {"value": t["value"], "unit": "channel*Msamples/s", "cpu_baseline": t,
                     "e2e": {"value": t["value"], "unit": "channel*Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ============================================================================ B200 arm
def pin_rank_to_cores(local_rank, world):
    """every rank on its own slice of the host cores: 8 Python ranks issuing ~30 launches per 0.5 ms step must not share cores"""
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // world)
        mine = cores[local_rank * per: (local_rank + 1) * per] or cores
        os.sched_setaffinity(0, mine)
        return len(mine)
    except Exception:
        return None


def sustained_loop(svc, issue, seconds, min_iters=8):
    """back-to-back `issue()` calls (asynchronous launches only) for >= `seconds` of DEVICE time; one CUDA-event pair brackets each
    chunk of iterations and the chunks grow, so the device queue only drains at the few chunk boundaries"""
    iters, total_ms = 0, 0.0
    chunk = min_iters
    wall0 = time.perf_counter()
    while total_ms < seconds * 1e3:
        svc.timer_start()
        for _ in range(chunk):
            issue()
        total_ms += svc.timer_stop()
        iters += chunk
        chunk = int(min(4096, max(chunk, chunk * 1.5)))
        if time.perf_counter() - wall0 > 4 * seconds + 5:
            break
    return iters, total_ms, wall0, time.perf_counter()


def parity_check_c2(res, stream_np, codes):
    """the bench's own acquisition results against the CPU oracle (tests/oracle_lib.OracleAcq), all 32 handles, untimed"""
    from concurrent.futures import ThreadPoolExecutor

    import oracle_lib as ol

    cfg = dict(sampled_ms=4, ms_per_code=4, samples_per_chip=8, doppler_max=DOPPLER_MAX, doppler_step=DOPPLER_STEP, fs_in=FS, samples_per_ms=8000.0,
               samples_per_code=float(N_FFT), threshold=ACQ_THRESHOLD, cshort=1)
    block = np.ascontiguousarray(stream_np[: 2 * N_FFT])

    def one(k):
        out = []
        for f64 in (0, 1):
            o = ol.OracleAcq(fft_f64=f64, **cfg)
            o.set_local_code(codes[k])
            r = o.run(block, N_FFT)
            out.append((r.code_phase_index, r.doppler_bin, r.decision, r.test_statistic, r.input_power))
            o.close()
        return out

    with ThreadPoolExecutor(min(16, os.cpu_count() or 2)) as ex:
        ora = list(ex.map(one, range(len(codes))))
    equal = cond = 0
    worst = 0.0
    for r, o in zip(res, ora):
        if o[0][:2] != o[1][:2]:
            continue  # float32 / float64 oracles disagree on the cell: ill-conditioned noise maximum, excluded
        cond += 1
        if (r.code_phase_index, r.doppler_bin, r.decision) == o[0][:3]:
            equal += 1
        worst = max(worst, abs(r.test_statistic - o[0][3]) / o[0][3], abs(r.input_power - o[0][4]) / o[0][4])
    return {"handles": len(res), "conditioned": cond, "cell_bin_decision_equal": equal, "worst_rel_statistic": worst, "ok": bool(equal == cond and worst <= 1e-4)}


def config_results(svc, gb, steps):
    """C1 / C3 / C5 on their own shapes (device-resident, L2 flushed; acquisition also through host buffers)."""
    import synth

    out = {}
    rng = np.random.default_rng(7)

    def acq_case(fs, n, kind, rate, prns, step, item, sampled_ms, spc, spms, cfar, scene_sats, cfar_threshold=1e-3):
        x = synth.make_iq(fs, n + 64, scene_sats, 1234)
        host = x if item == gb.ITEM_32FC else synth.quantize(x, "16ic")
        svc.push(2, host, item, 0)
        acqs = []
        for p in prns:
            a = gb.Acquisition(svc, sampled_ms=sampled_ms, ms_per_code=sampled_ms, samples_per_chip=spc, doppler_max=5000, doppler_step=step, fs_in=fs,
                               samples_per_ms=spms, samples_per_code=float(n), threshold=cfar_threshold if cfar else 2.0, use_CFAR_algorithm_flag=int(cfar), item_type=item)
            a.set_local_code(synth.sampled_code(synth.code_table(kind, p), rate, fs, n))
            acqs.append(a)
        h = gb.acq_handles(acqs)
        cells = len(prns) * acqs[0].num_bins * n
        ms, e2e = [], []
        for it in range(3 + steps):
            svc.flush_l2()
            svc.timer_start()
            gb.acq_launch_batch(acqs, h, 2, 0)
            t = svc.timer_stop()
            res = gb.acq_collect_batch(acqs, h)
            svc.flush_l2()
            svc.timer_start()
            svc.push(2, host[: (n if item == gb.ITEM_32FC else 2 * n)], item, 0)
            res = gb.acq_run_batch(acqs, 2, 0)
            te = svc.timer_stop()
            if it >= 3:
                ms.append(t)
                e2e.append(te)
        found = sum(r.decision == 1 for r in res)
        for a in acqs:
            a.close()
        return {"cells_per_step": cells, "ms_per_step": float(np.mean(ms)), "value": cells / (np.mean(ms) * 1e-3), "e2e": cells / (np.mean(e2e) * 1e-3),
                "unit": "grid cells/s", "positives": int(found), "present": len(scene_sats)}

    def trk_case(fs, n_ch, n_epoch, epoch, kind, dkind, taps, item, table_len, carrier):
        n_items = min(1 << 23, n_epoch * epoch + 40000)
        n_epoch = min(n_epoch, (n_items - 40000) // epoch)
        x = (rng.standard_normal(n_items, np.float32) + 1j * rng.standard_normal(n_items, np.float32)).astype(np.complex64)
        host = x if item == gb.ITEM_32FC else synth.quantize(x * 8.0, "16ic")
        svc.push(3, host, item, 0)
        chans = []
        for c in range(n_ch):
            m = gb.Multicorrelator(svc, len(taps), int(table_len), item_type=item)
            m.set_local_code_and_taps(synth.code_table(kind, 1 + c), taps)
            if dkind:
                m.set_data_code(synth.code_table(dkind, 1 + c), len(taps) // 2)
            chans.append(m)
        rl = trk_requests(0, n_ch, n_epoch, epoch, fs, table_len, carrier)
        reqs = gb.make_reqs((chans[c], 3, first, rc, ps, 0.0, rcode, cs, 0.0, epoch) for (c, first, rc, ps, rcode, cs) in rl)
        batch = gb.TrkBatch(svc, reqs)
        ms = []
        for it in range(3 + steps):
            svc.flush_l2()
            svc.timer_start()
            batch.launch()
            t = svc.timer_stop()
            if it >= 3:
                ms.append(t)
        chs = len(rl) * epoch
        batch.close()
        for m in chans:
            m.close()
        flops = 12 + 4 * (len(taps) + (1 if dkind else 0))
        return {"channel_samples_per_step": chs, "ms_per_step": float(np.mean(ms)), "value": chs / (np.mean(ms) * 1e-3) / 1e6, "unit": "channel*Msamples/s",
                "flops_per_channel_sample": flops, "tflops": flops * chs / (np.mean(ms) * 1e-3) / 1e12}

    gps = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=50 - k, doppler_hz=-4375 + 1250 * k, delay_samples=137 + 487 * k) for k, p in enumerate([1, 3, 6, 9, 12, 17, 22, 28])]
    out["C1"] = {"workload": "GPS L1 C/A, 4 Msps gr_complex, 8 channels: acquisition 8 x 40 bins x 4000; tracking 8 ch x 2000 epochs x 4000 samples, 3 taps",
                 "acquisition": acq_case(4_000_000, 4000, "gps_l1_ca", 1.023e6, [1, 3, 6, 9, 12, 17, 22, 28], 250, gb.ITEM_32FC, 1, 4, 4000.0, True, gps),
                 "tracking": trk_case(4e6, 8, 2000, 4000, "gps_l1_ca", None, np.array([-0.5, 0.0, 0.5], np.float32), gb.ITEM_32FC, 1023.0, 1575.42e6)}
    l5 = [dict(kind="gps_l5i", prn=p, cn0_dbhz=53 - k, doppler_hz=500 + 700 * k, delay_samples=1111 + 3571 * k) for k, p in enumerate([1, 3, 6, 10, 24, 25])]
    out["C3"] = {"workload": "GPS L5 / Galileo E5a 10230-chip codes, 25 Msps gr_complex: acquisition 12 x 40 bins x 25000 (first/second-peak statistic); tracking 12 ch x 320 epochs x 25000 samples, 3 taps + data-prompt tap",
                 "acquisition": acq_case(25_000_000, 25000, "gps_l5i", 10.23e6, [1, 3, 6, 10, 24, 25, 2, 4, 5, 7, 8, 9], 250, gb.ITEM_32FC, 1, 3, 25000.0, False, l5),
                 "tracking": trk_case(25e6, 12, 320, 25000, "gps_l5q", "gps_l5i", np.array([-0.5, 0.0, 0.5], np.float32), gb.ITEM_32FC, 10230.0, 1176.45e6)}
    e1 = [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=p, cn0_dbhz=50 - 2 * k, doppler_hz=-3300 + 2900 * k, delay_samples=12345 + 61111 * k) for k, p in enumerate([2, 12, 25])]
    out["C5"] = {"workload": "cold start at 50 Msps gr_complex: Galileo E1 acquisition 8 PRN x 80 bins x 200000 (of the 36-PRN search; --mode shard runs all 68 PRNs); tracking 8 E1 pilot channels x 40 epochs x 200000 samples, 5 taps + data-prompt tap",
                 "acquisition": acq_case(50_000_000, 200000, "gal_e1b_sinboc", 2 * 1.023e6, [2, 12, 25, 1, 3, 4, 5, 6], 125, gb.ITEM_32FC, 4, 49, 50000.0, True, e1,
                                         cfar_threshold=2e-4),  # E1B alone carries half the power: statistic ~ CN0 / (2 fs) = 1e-3 at 50 dBHz; noise-only maximum ~ ln(cells) / N = 8e-5
                 "tracking": trk_case(50e6, 8, 40, 200000, "gal_e1c_sinboc", "gal_e1b_sinboc", TAPS, gb.ITEM_32FC, 8184.0, 1575.42e6)}
    return out


def run_b200(args, rank, local_rank, world):
    import torch

    import gnss_b200 as gb

    dist = None
    cores_mine = None
    if world > 1:
        import torch.distributed as dist

        cores_mine = pin_rank_to_cores(local_rank, world)
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank
    # ---- the sample stream: rank 0 synthesises it; NCCL broadcasts the packed int16 items to every GPU
    if rank == 0:
        stream_np, sats = build_stream(20260002)
    else:
        stream_np = np.empty(2 * STREAM_ITEMS, np.int16)
    host_stream = torch.empty(2 * STREAM_ITEMS, dtype=torch.int16).pin_memory()
    if world > 1:
        d_stream = torch.from_numpy(stream_np).to("cuda:%d" % dev) if rank == 0 else torch.empty(2 * STREAM_ITEMS, dtype=torch.int16, device="cuda:%d" % dev)
        dist.broadcast(d_stream.view(torch.uint8), src=0)  # NCCL has no int16: ship the packed items as bytes
        host_stream.copy_(d_stream.cpu())
        del d_stream
    else:
        host_stream.copy_(torch.from_numpy(stream_np))
    host_ptr = host_stream.data_ptr()

    svc = gb.Service(device=dev, ring_capacity=1 << 24, ring_slack=1 << 20)
    tuned = svc.tune_from_env()  # GB200_TUNE (profiling sweeps only; recorded in the line when set)
    fp32_peak = svc.measure_fp32_tflops()
    hbm_peak, peak_src = measured_peaks()

    # ---- acquisition handles (one per PRN of this rank's shard; weak scaling: every rank searches 32 codes)
    codes = acq_codes()
    acqs = []
    for p in range(N_PRN):
        a = gb.Acquisition(svc, sampled_ms=4, ms_per_code=4, samples_per_chip=8, doppler_max=DOPPLER_MAX, doppler_step=DOPPLER_STEP,
                           fs_in=FS, samples_per_ms=8000.0, samples_per_code=float(N_FFT), threshold=ACQ_THRESHOLD, item_type=gb.ITEM_16IC)
        a.set_local_code(np.roll(codes[p], 131 * rank))
        acqs.append(a)
    handles = gb.acq_handles(acqs)

    # ---- tracking channels + staged epoch requests
    import synth

    chans = []
    for c in range(N_CH):
        prn = 1 + (c % 32)
        m = gb.Multicorrelator(svc, 5, 8184, item_type=gb.ITEM_16IC)
        m.set_local_code_and_taps(synth.code_table("gal_e1c_sinboc", prn), TAPS)
        m.set_data_code(synth.code_table("gal_e1b_sinboc", prn), 2)
        chans.append(m)
    req_list = trk_requests(rank)
    svc.push_ptr(0, host_ptr, STREAM_ITEMS, gb.ITEM_16IC, 0)
    reqs = gb.make_reqs((chans[c], 0, first, rc, ps, 0.0, rcode, cs, 0.0, EPOCH) for (c, first, rc, ps, rcode, cs) in req_list)
    batch = gb.TrkBatch(svc, reqs)
    svc.synchronize()

    ACQ_BLOCK_FIRST = 0
    acq_cells = N_PRN * N_BINS * N_FFT
    trk_chsamples = len(req_list) * EPOCH

    def barrier():
        svc.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step(timed):
        """inputs resident in HBM; returns (acq_ms, trk_ms) from CUDA events on the service stream"""
        svc.flush_l2()
        svc.timer_start()
        gb.acq_launch_batch(acqs, handles, 0, ACQ_BLOCK_FIRST)
        a_ms = svc.timer_stop()
        res = gb.acq_collect_batch(acqs, handles)
        svc.flush_l2()
        svc.timer_start()
        batch.launch()
        t_ms = svc.timer_stop()
        return a_ms, t_ms, res

    for _ in range(max(3, args.warmup)):
        a_ms, t_ms, res = device_step(False)
    launches0 = svc.launch_count()
    barrier()
    sampler = ClockSampler(dev) if rank == 0 else None
    if sampler:
        sampler.start()
    t_burst0 = time.perf_counter()
    acq_ms, trk_ms = [], []
    for i in range(args.steps):
        # per-launch CUDA events (two records around every kernel launch, which also stop consecutive kernels from
        # overlapping their tails) cost ~20 % of an acquisition step, so only the LAST timed step carries them
        svc.kernel_timing(i == args.steps - 1)
        a_ms, t_ms, res = device_step(True)
        acq_ms.append(a_ms)
        trk_ms.append(t_ms)
    barrier()
    kstats = {k: svc.kernel_stats(v) for k, v in gb.KCLASS.items()}
    svc.kernel_timing(False)
    launches = svc.launch_count() - launches0
    out_trk = batch.fetch()

    # ---- end to end through the C ABI with HOST buffers: H2D of the step's samples + launch + D2H of the results
    e2e_acq_ms, e2e_trk_ms = [], []
    # tracking: the second of samples is pushed in E2E_BLOCKS blocks of whole epochs
    per_blk = N_EPOCH // E2E_BLOCKS
    blk_edges = [0] + [min(STREAM_ITEMS, max(r[1] for r in req_list) - (N_EPOCH - (k + 1) * per_blk) * EPOCH + EPOCH) for k in range(E2E_BLOCKS)]
    blk_edges[-1] = STREAM_ITEMS
    for it in range(max(2, args.warmup // 2) + args.steps):
        svc.flush_l2()
        svc.timer_start()
        svc.push_ptr(1, host_ptr, N_FFT, gb.ITEM_16IC, 0)       # the 4 ms block acquisition consumes
        res_e2e = gb.acq_run_batch(acqs, 1, 0)                     # launch + D2H of 32 result records
        ea = svc.timer_stop()
        svc.flush_l2()
        svc.timer_start()
        # 1 s of samples for the 248 epochs, handed over in E2E_BLOCKS blocks the way a front-end delivers them (asynchronous
        # copies on the ingest stream); ONE correlate call for all epochs: the service launches each block's epochs as soon
        # as that block's push has landed, so the H2D copy of block k+1 overlaps the correlation of block k
        for k in range(E2E_BLOCKS):
            svc.push_ptr(0, host_ptr + 4 * blk_edges[k], blk_edges[k + 1] - blk_edges[k], gb.ITEM_16IC, blk_edges[k])
        out_e2e = gb.correlate_batch(svc, reqs)                    # stage requests + launches + D2H of correlator outputs
        et = svc.timer_stop()
        if it >= max(2, args.warmup // 2):
            e2e_acq_ms.append(ea)
            e2e_trk_ms.append(et)
    t_e2e1 = time.perf_counter()

    # ---- sustained: >= args.sustain seconds per path of back-to-back steps, no host synchronisation or L2 flush inside
    sustained = None
    if args.sustain > 0:
        barrier()

        def issue_acq():
            gb.acq_launch_batch(acqs, handles, 0, ACQ_BLOCK_FIRST)
            gb.acq_discard_batch(acqs, handles)

        it_a, ms_a, ta0, ta1 = sustained_loop(svc, issue_acq, args.sustain)
        svc.synchronize()
        it_t, ms_t, tt0, tt1 = sustained_loop(svc, batch.launch, args.sustain)
        svc.synchronize()
        sustained = {"seconds_acq": ms_a * 1e-3, "seconds_trk": ms_t * 1e-3, "acq_steps": it_a, "trk_steps": it_t,
                     "acq_ms_per_step": ms_a / it_a, "trk_ms_per_step": ms_t / it_t, "t": (ta0, ta1, tt0, tt1)}
    if sampler:
        sampler.stop_flag = True
        sampler.join(timeout=2)

    # ---- sanity: the bench's own results are right (not timed): satellites found where they were put
    found = {(1 + 3 * k): (res[3 * k].code_phase_index, res[3 * k].doppler_hz, res[3 * k].decision) for k in range(10)} if rank == 0 else {}
    assert np.isfinite(out_trk[:, :6]).all() and np.abs(out_trk[:, 2]).max() > 0
    assert np.allclose(out_trk[:, :6], out_e2e[:, :6], rtol=1e-5, atol=1e-2)

    vals = [sum(acq_ms), sum(trk_ms), sum(e2e_acq_ms), sum(e2e_trk_ms)]
    if sustained:
        vals += [sustained["acq_ms_per_step"], sustained["trk_ms_per_step"]]
    t_acq = torch.tensor(vals, dtype=torch.float64, device="cuda:%d" % dev)
    if dist is not None:
        dist.all_reduce(t_acq, op=dist.ReduceOp.MAX)
    vals = [float(v) for v in t_acq.cpu()]
    tot_acq_ms, tot_trk_ms, tot_e2e_acq_ms, tot_e2e_trk_ms = vals[:4]
    steps = args.steps
    acq_value = world * acq_cells * steps / (tot_acq_ms * 1e-3)
    trk_value = world * trk_chsamples * steps / (tot_trk_ms * 1e-3) / 1e6
    e2e_acq = world * acq_cells * steps / (tot_e2e_acq_ms * 1e-3)
    e2e_trk = world * trk_chsamples * steps / (tot_e2e_trk_ms * 1e-3) / 1e6

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- rooflines (per launch, live CUDA-event durations of the kernel classes)
    inv_row_ms, inv_row_n = kstats["acq_inv_row"]
    inv_col_ms, inv_col_n = kstats["acq_inv_col"]
    n_pairs = max(1, max(inv_col_n, inv_row_n))
    pair_ms = (inv_row_ms + inv_col_ms) / n_pairs or float('nan')
    cells_per_launch = acq_cells / n_pairs
    acq_gbs = 16.0 * cells_per_launch / (pair_ms * 1e-3) / 1e9
    trk_ms_k, trk_n = kstats["trk"]
    flops_per_chs = 12 + 4 * 6
    trk_tflops = flops_per_chs * trk_chsamples / ((trk_ms_k or float('nan')) * 1e-3) / 1e12
    traffic = traffic_constants()
    flop_per_cell = 10.0 * np.log2(N_FFT) + 16.0
    sus_line = None
    if sustained:
        s_acq = world * acq_cells / (vals[4] * 1e-3)
        s_trk = world * trk_chsamples / (vals[5] * 1e-3) / 1e6
        ta0, ta1, tt0, tt1 = sustained.pop("t")
        sus_line = {"note": "back-to-back steps, asynchronous launches only (acquisition: launch + discard, i.e. no result read-back), no L2 flush, one CUDA-event pair per chunk of steps",
                    "acq": {"value": s_acq, "unit": "grid cells/s", "ms_per_step": vals[4], "seconds": sustained["seconds_acq"], "steps": sustained["acq_steps"],
                            "roofline_frac_hbm16": s_acq / world * 16.0 / 1e9 / hbm_peak, "fp32_frac": s_acq / world * flop_per_cell / 1e12 / fp32_peak if fp32_peak else None,
                            "clocks": sampler.summary(ta0, ta1)},
                    "trk": {"value": s_trk, "unit": "channel*Msamples/s", "ms_per_step": vals[5], "seconds": sustained["seconds_trk"], "steps": sustained["trk_steps"],
                            "roofline_frac_fp32": s_trk / world * 1e6 * flops_per_chs / 1e12 / fp32_peak if fp32_peak else None,
                            "clocks": sampler.summary(tt0, tt1)}}
    parity = None
    cfg_res = None
    if os.environ.get("GB_BENCH_SKIP_CPU"):  # profiling runs only (ncu): never used for a reported line
        cpu_acq = cpu_trk = {"value": None, "skipped": "GB_BENCH_SKIP_CPU set (profiling run)"}
    else:
        try:
            parity = parity_check_c2(res, stream_np, [np.roll(c, 131 * rank) for c in codes])
        except Exception as e:  # the oracle is the checker, not the product: report, do not hide the measurement
            parity = {"ok": None, "error": repr(e)}
        if world == 1 and not args.no_configs:
            cfg_res = config_results(svc, gb, max(3, min(args.steps, 10)))
        cpu_acq = cpu_acquisition_baseline(stream_np, codes, 8.0)
        cpu_trk = cpu_tracking_baseline(8.0)
    line = {
        "metric": METRIC, "value": acq_value, "unit": "grid cells/s", "n_gpus": world, "steps": steps, "warmup": max(3, args.warmup),
        "ms_per_step": (tot_acq_ms + tot_trk_ms) / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "l2": "flushed (256 MB write) before every timed region",
                   "kernel_events": "per-launch CUDA events on the library stream during the last timed step only", "acq_ms_per_step": tot_acq_ms / steps, "acq_ms_steps": [round(v, 4) for v in acq_ms], "trk_ms_per_step": tot_trk_ms / steps,
                   "host_cores_per_rank": cores_mine, "tuning": tuned or None},
        "e2e": {"value": e2e_acq, "unit": "grid cells/s", "h2d_bytes_per_step": N_FFT * 4, "d2h_bytes_per_step": N_PRN * 56,
                "ms_per_step": tot_e2e_acq_ms / steps},
        "tracking_value": trk_value, "tracking_unit": "channel*Msamples/s", "tracking_e2e": e2e_trk,
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": acq_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": acq_gbs / hbm_peak, "traffic": traffic.get("acq_pair_dram_bytes"),
                     "traffic_note": "dram__bytes_read+write of one inverse launch pair from ncu --set full, per launch like `achieved`: " + str(traffic.get("source")) +
                                     "; ncu flushes caches between kernels, so the column kernel's read of the row buffer is DRAM there and an L2 hit in the pipeline; algorithmic bytes of the pair: %.4g" % (16.0 * cells_per_launch),
                     "kernel": "inverse side of one chunk of (PRN, bin) transforms: row pass + column pass (16 algorithmic B per cell)",
                     "peak_source": peak_src, "per_launch_ms": pair_ms, "cells_per_launch": cells_per_launch,
                     "step_wall_frac": acq_value / world * 16.0 / 1e9 / hbm_peak,
                     "fp32_view": {"flop_per_cell": flop_per_cell, "achieved_tflops": acq_value / world * flop_per_cell / 1e12, "peak_tflops": fp32_peak,
                                   "frac": acq_value / world * flop_per_cell / 1e12 / fp32_peak if fp32_peak else None},
                     "kernel_ms": {k: {"total_ms": v[0], "launches": v[1]} for k, v in kstats.items()}},
        "cpu_baseline": cpu_acq,
        "clocks": sampler.summary(t_burst0, t_e2e1),
        "tracking": {
            "metric": "tracking channel*Msamples/s", "value": trk_value, "unit": "channel*Msamples/s",
            "e2e": {"value": e2e_trk, "unit": "channel*Msamples/s", "h2d_bytes_per_step": STREAM_ITEMS * 4 + len(req_list) * 120,
                    "d2h_bytes_per_step": len(req_list) * 72, "ms_per_step": tot_e2e_trk_ms / steps,
                    "blocks": E2E_BLOCKS, "note": "samples pushed in %d blocks on the ingest stream; one correlate call, each block's epochs launch when its push has landed" % E2E_BLOCKS},
            "roofline": {"bound": "fp32", "achieved": trk_tflops, "peak": fp32_peak, "unit": "TFLOP/s", "frac": trk_tflops / fp32_peak if fp32_peak else None,
                         "traffic": traffic.get("trk_dram_bytes"), "traffic_note": "dram__bytes_read+write of the launch under ncu --set full (" + str(traffic.get("source")) + "): the 32 MB sample stream once + request records; the 32 channels re-read it from L2",
                         "kernel": "gb_trk_kernel<16IC,5,data,prompt=2>", "flops_per_channel_sample": flops_per_chs,
                         "peak_source": "FMA micro-benchmark run in this process (gb200_service_measure_fp32)"},
            "cpu_baseline": cpu_trk,
        },
        "sustained": sus_line,
        "config_results": cfg_res,
        "parity": parity,
        "acq_found": {str(k): list(map(int, v)) for k, v in found.items()},
    }
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="weak", choices=["weak", "shard"])
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of back-to-back steps per path (0: skip)")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C3 / C5 lines")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.mode == "shard":
        import bench_shard

        bench_shard.run(args, rank, local_rank, world)
        return
    run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
