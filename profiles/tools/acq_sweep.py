"""Profiling helper (not part of the product or the bench): sustained C2 acquisition step time under tuning switches.
usage: python profiles/tools/acq_sweep.py "acq_l2_mb=48,acq_lanes=2" "acq_l2_mb=200,acq_lanes=1" ...   (one service per spec)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gnss-sdr-1_b200")]
import gnss_b200 as gb  # noqa: E402

N, PRNS = 32000, 32


def run(spec, seconds=1.0):
    s = gb.Service(device=0, ring_capacity=1 << 22, ring_slack=1 << 20)
    for kv in [x for x in spec.split(",") if x]:
        k, v = kv.split("=")
        s.set_tuning(k, int(v))
    rng = np.random.default_rng(1)
    x = rng.integers(-200, 200, 2 * (N + 64)).astype(np.int16)
    s.push(0, x, gb.ITEM_16IC, 0)
    acqs = []
    for p in range(PRNS):
        a = gb.Acquisition(s, sampled_ms=4, ms_per_code=4, samples_per_chip=8, doppler_max=10000, doppler_step=250, fs_in=8000000, samples_per_ms=8000.0,
                           samples_per_code=float(N), threshold=1e9, item_type=gb.ITEM_16IC)
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
    for a in acqs:
        a.close()
    s.close()


if __name__ == "__main__":
    for spec in (sys.argv[1:] or [""]):
        run(spec)
