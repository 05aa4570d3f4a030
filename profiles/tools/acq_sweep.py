"""Profiling helper (not part of the product or the bench): sustained C2 acquisition step time under tuning switches.
usage: python profiles/tools/acq_sweep.py "acq_l2_mb=48,acq_lanes=2" "acq_l2_mb=200,acq_lanes=1" ...   (one service per spec)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gnss-sdr-1_b200")]
import gnss_b200 as gb  # noqa: E402

# ACQ_SWEEP_CFG=C2 (default) | C3 | C5: FFT length, handles, item type, Doppler span, CFAR flag of the BASELINE config
CFG = {"C2": (32000, 32, "16ic", 10000, 8000000, 4, 1), "C3": (25000, 12, "32fc", 5000, 25000000, 1, 0), "C5": (200000, 8, "32fc", 10000, 50000000, 4, 1)}[os.environ.get("ACQ_SWEEP_CFG", "C2")]
if os.environ.get("ACQ_SWEEP_N"):  # any other FFT length: 8 handles, 1 ms at N ksps, +-5 kHz / 250 Hz, CFAR, gr_complex
    n_ = int(os.environ["ACQ_SWEEP_N"])
    CFG = (n_, 8, "32fc", 5000, n_ * 1000, 1, 1)
N, PRNS = CFG[0], CFG[1]


def run(spec, seconds=1.0):
    s = gb.Service(device=0, ring_capacity=1 << 22, ring_slack=1 << 20)
    for kv in [x for x in spec.split(",") if x]:
        k, v = kv.split("=")
        s.set_tuning(k, int(v))
    rng = np.random.default_rng(1)
    item = gb.ITEM_16IC if CFG[2] == "16ic" else gb.ITEM_32FC
    if CFG[2] == "16ic":
        x = rng.integers(-200, 200, 2 * (N + 64)).astype(np.int16)
    else:
        x = (rng.standard_normal(N + 64) + 1j * rng.standard_normal(N + 64)).astype(np.complex64)
    s.push(0, x, item, 0)
    acqs = []
    for p in range(PRNS):
        a = gb.Acquisition(s, sampled_ms=CFG[5], ms_per_code=CFG[5], samples_per_chip=8, doppler_max=CFG[3], doppler_step=250, fs_in=CFG[4], samples_per_ms=N / CFG[5],
                           samples_per_code=float(N), threshold=1e9, use_CFAR_algorithm_flag=CFG[6], item_type=item)
        a.set_local_code((rng.integers(0, 2, N) * 2 - 1).astype(np.complex64))
        acqs.append(a)
    h = gb.acq_handles(acqs)
    for _ in range(5):
        gb.acq_launch_batch(acqs, h, 0, 0)
        gb.acq_discard_batch(acqs, h)
    s.synchronize()
    iters, total = 0, 0.0
    chunk = 16
    while total < seconds * 1e3:
        s.timer_start()
        for _ in range(chunk):
            gb.acq_launch_batch(acqs, h, 0, 0)
            gb.acq_discard_batch(acqs, h)
        total += s.timer_stop()
        iters += chunk
        chunk = min(1024, chunk * 2)
    ms = total / iters
    print("%-60s %.4f ms/step  %.4g cells/s  (bins %d)" % (spec or "(defaults)", ms, PRNS * acqs[0].num_bins * N / (ms * 1e-3), acqs[0].num_bins), flush=True)
    if os.environ.get("ACQ_SWEEP_KERNELS"):
        s.kernel_timing(True)
        gb.acq_launch_batch(acqs, h, 0, 0)
        gb.acq_collect_batch(acqs, h)
        print("    " + "  ".join("%s %.4f ms / %d" % ((name,) + tuple(s.kernel_stats(k))) for k, name in enumerate(("trk", "fwd_col", "fwd_row", "inv_row", "inv_col")) if k), flush=True)
    for a in acqs:
        a.close()
    s.close()


if __name__ == "__main__":
    for spec in (sys.argv[1:] or [""]):
        run(spec)
