"""World-size-2 (gloo, CPU) tests of the multi-GPU host logic in gnss-sdr-1_b200/shard.py: unit partitioning, the
Doppler-bin-range trick through freq_offset_hz, record gathering and the tie-break-preserving merge.  The oracle stands in for
the device compute here (no GPU in this container); the same functions drive bench.py under torchrun on GPUs."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [HERE, os.path.join(HERE, "..", "gnss-sdr-1_b200")]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _signal(n, prns, seed):
    import synth

    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    t = np.arange(n) / 4e6
    truth = {}
    for k, prn in enumerate(prns):
        code = synth.sampled_code(synth.code_table("gps_l1_ca", prn), 1.023e6, 4e6, n)
        delay, dop = 300 + 577 * k, -3750.0 + 1500.0 * k
        x += (0.6 * np.roll(code, delay) * np.exp(2j * np.pi * dop * t)).astype(np.complex64)
        truth[prn] = (delay, dop)
    return x, truth


def _worker(rank, world, port, n_prn, q):
    import torch.distributed as dist

    import oracle_lib as ol
    import shard
    import synth

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prns = [1, 6, 11, 17][:n_prn]
        n = 4000
        x, _ = _signal(n, prns, 77)
        dmax, step = 5000, 250
        units = shard.acquisition_units(n_prn, dmax, step, world)
        recs = []
        for u in shard.partition(len(units), rank, world):
            unit = units[u]
            o = ol.OracleAcq(doppler_max=unit.doppler_max, doppler_step=unit.doppler_step, freq_offset_hz=unit.freq_offset_hz, threshold=0.0008)
            assert o.num_bins == unit.bin_hi - unit.bin_lo
            o.set_local_code(synth.sampled_code(synth.code_table("gps_l1_ca", prns[unit.prn_slot]), 1.023e6, 4e6, n).astype(np.complex64))
            r = o.run(x)
            recs.append(shard.acq_record(u, unit, r.code_phase_index, r.doppler_bin, r.doppler_hz, r.peak_mag, r.input_power))
            o.close()
        allrecs = shard.gather_records(recs, dist, dst=0)
        # tracking: channels dealt round-robin, outputs gathered in channel order
        n_ch, epochs, taps = 5, 3, 3
        mine = shard.partition(n_ch, rank, world)
        outs = np.zeros((len(mine), epochs, taps), np.complex64)
        code = synth.code_table("gps_l1_ca", 1)
        for k, c in enumerate(mine):
            for e in range(epochs):
                outs[k, e] = ol.multicorrelator_real("oracle", x[:2000], code, [-0.5, 0.0, 0.5], 0.1 * c, 0.01 * (e + 1), 0.25 * c, 0.2557)
        full = shard.gather_channel_outputs(mine, outs, n_ch, dist, dst=0)
        if rank == 0:
            merged = shard.merge_acq_records(allrecs, n_prn, n, threshold=0.0008)
            ref = []
            for prn in prns:
                o = ol.OracleAcq(doppler_max=dmax, doppler_step=step, threshold=0.0008)
                o.set_local_code(synth.sampled_code(synth.code_table("gps_l1_ca", prn), 1.023e6, 4e6, n).astype(np.complex64))
                r = o.run(x)
                ref.append((r.code_phase_index, r.doppler_bin, r.doppler_hz, r.peak_mag, r.test_statistic, r.decision))
                o.close()
            want = np.zeros((n_ch, epochs, taps), np.complex64)
            for c in range(n_ch):
                for e in range(epochs):
                    want[c, e] = ol.multicorrelator_real("oracle", x[:2000], code, [-0.5, 0.0, 0.5], 0.1 * c, 0.01 * (e + 1), 0.25 * c, 0.2557)
            q.put((len(allrecs), len(units), merged, ref, bool(np.array_equal(full, want))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_prn", [4, 1])
def test_world2_sharded_search_equals_unsharded(n_prn):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_prn, q)) for r in range(2)]
    for p in procs:
        p.start()
    import queue
    import time

    got, t0 = None, time.time()
    while got is None and time.time() - t0 < 240:
        try:
            got = q.get(timeout=2)
        except queue.Empty:
            if any(p.exitcode not in (None, 0) for p in procs):
                break
    if got is None:
        for p in procs:
            if p.is_alive():
                p.terminate()
        pytest.fail("a rank died or timed out")
    n_recs, n_units, merged, ref, trk_ok = got
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert n_recs == n_units == (4 if n_prn == 4 else 2)
    assert trk_ok
    for m, (idx, b, hz, mag, stat, dec) in zip(merged, ref):
        assert (m["code_phase_index"], m["doppler_bin"], m["doppler_hz"], m["decision"]) == (idx, b, hz, dec)
        assert m["peak_mag"] == pytest.approx(mag, rel=1e-6) and m["test_statistic"] == pytest.approx(stat, rel=1e-5)
        assert dec == 1


def test_units_cover_grid_exactly():
    import shard

    for n_prn, world, dmax, step in [(32, 8, 5000, 125), (1, 8, 5000, 250), (3, 8, 5000, 250), (2, 4, 10000, 500), (1, 2, 5000, 333)]:
        units = shard.acquisition_units(n_prn, dmax, step, world)
        nb = shard.num_doppler_bins(dmax, step)
        for p in range(n_prn):
            mine = sorted((u.bin_lo, u.bin_hi) for u in units if u.prn_slot == p)
            assert mine[0][0] == 0 and mine[-1][1] == nb
            assert all(a[1] == b[0] for a, b in zip(mine, mine[1:]))
        for u in units:
            # the sub-grid's own bin i maps to the reference grid's wipe-off frequency
            assert u.freq_offset_hz + (-u.doppler_max) == -dmax + step * u.bin_lo
            assert shard.num_doppler_bins(u.doppler_max, u.doppler_step) == u.bin_hi - u.bin_lo
        owners = [len(shard.partition(len(units), r, world)) for r in range(world)]
        assert sum(owners) == len(units) and max(owners) - min(owners) <= 1
