"""Profiling helper: where does a C3 acquisition step (12 handles x 40 bins x 25000 points, first/second-peak statistic) spend its time?
usage: python profiles/tools/acq_c3_probe.py [cfar]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "gnss-sdr-1_b200")]
import gnss_b200 as gb  # noqa: E402

cfar = len(sys.argv) > 1 and sys.argv[1] == "cfar"
N, H = 25000, 12
s = gb.Service(device=0, ring_capacity=1 << 22, ring_slack=1 << 20)
rng = np.random.default_rng(1)
x = (rng.standard_normal(N + 64) + 1j * rng.standard_normal(N + 64)).astype(np.complex64)
s.push(0, x, gb.ITEM_32FC, 0)
acqs = []
for p in range(H):
    a = gb.Acquisition(s, sampled_ms=1, ms_per_code=1, samples_per_chip=3, doppler_max=5000, doppler_step=250, fs_in=25000000, samples_per_ms=25000.0,
                       samples_per_code=float(N), threshold=2.0, use_CFAR_algorithm_flag=int(cfar), item_type=gb.ITEM_32FC)
    a.set_local_code((rng.integers(0, 2, N) * 2 - 1).astype(np.complex64))
    acqs.append(a)
h = gb.acq_handles(acqs)
for _ in range(5):
    gb.acq_launch_batch(acqs, h, 0, 0)
    gb.acq_collect_batch(acqs, h)
for mode in ("launch+collect", "launch+discard"):
    s.synchronize()
    t0 = time.perf_counter()
    s.timer_start()
    for _ in range(200):
        gb.acq_launch_batch(acqs, h, 0, 0)
        (gb.acq_collect_batch if mode == "launch+collect" else gb.acq_discard_batch)(acqs, h)
    dev = s.timer_stop() / 200
    wall = (time.perf_counter() - t0) / 200 * 1e3
    print("%s  %s: device %.4f ms/step, host wall %.4f ms/step" % ("CFAR" if cfar else "peak-ratio", mode, dev, wall))
s.kernel_timing(True)
n0 = s.launch_count()
gb.acq_launch_batch(acqs, h, 0, 0)
gb.acq_collect_batch(acqs, h)
print("launches in one step:", s.launch_count() - n0)
for k, name in enumerate(("trk", "fwd_col", "fwd_row", "inv_row", "inv_col")):
    ms, n = s.kernel_stats(k)
    print("  %-8s %.4f ms in %d launches" % (name, ms, n))
