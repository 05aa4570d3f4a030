"""bench.py --mode shard: strong scaling of ONE fixed search and ONE fixed channel set over the ranks.

  acquisition  C5 cold start at 50 Msps gr_complex: 32 GPS L1 C/A PRNs (N = 50000, 40 bins) + 36 Galileo E1 PRNs (N = 200000, 80 bins)
               = 640 M grid cells per step, dealt PRN by PRN over the ranks by gnss-sdr-1_b200/shard.py
  tracking     C4: 1024 channels on a 20 Msps interleaved-int8 stream, 40 ms per step: 256 GPS L1 (3 taps), 256 Galileo E1 pilot
               (5 taps + data-prompt tap), 256 BeiDou B1I (3 taps), 256 GLONASS L1 (3 taps, COMPLEX codes)  = 819.2 M channel.samples
Inside the timed region of every step, on every rank: the NCCL broadcast of the step's sample block from rank 0's device, the
device-to-ring push, the rank's own share through the C ABI (results back on the host), the NCCL all-gather of the 56-byte peak
records / the correlator outputs, and on rank 0 the merge with the reference's tie-break (magnitude desc, Doppler bin asc, index
asc).  Timed with CUDA events on the torch stream (the host waits for the service stream before it issues the gather, so the event
pair brackets all of it), max over ranks.  One JSON line, "scaling": "strong".
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def run(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    import gnss_b200 as gb
    import shard
    import synth
    import bench

    if world > 1:
        bench.pin_rank_to_cores(local_rank, world)
    torch.cuda.set_device(local_rank)
    backend_inited = False
    if world > 1 or os.environ.get("MASTER_ADDR"):
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        backend_inited = True
    dev = torch.device("cuda", local_rank)
    svc = gb.Service(device=local_rank, ring_capacity=1 << 22, ring_slack=1 << 20)
    svc.tune_from_env()

    # ------------------------------------------------------------------ acquisition: C5 cold start, 68 PRNs
    fs = 50_000_000
    n_gps, n_e1 = 50000, 200000
    sats = [dict(kind="gps_l1_ca", prn=p, cn0_dbhz=52 - k, doppler_hz=-4000 + 1100 * k, delay_samples=777 + 5003 * k) for k, p in enumerate([2, 5, 9, 13, 17, 21, 26, 30])]
    sats += [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=p, cn0_dbhz=52 - k, doppler_hz=-3800 + 1000 * k, delay_samples=12345 + 21111 * k)
             for k, p in enumerate([1, 4, 8, 11, 19, 24, 30, 36])]
    if rank == 0:
        x = synth.make_iq(fs, n_e1 + 1000, sats, 20260005)
        d_block = torch.from_numpy(x.view(np.float32).copy()).to(dev)
    else:
        d_block = torch.zeros(2 * (n_e1 + 1000), dtype=torch.float32, device=dev)
    prn_list = [("G", p) for p in range(1, 33)] + [("E", p) for p in range(1, 37)]
    units = shard.acquisition_units(len(prn_list), 5000, 250, world)   # 68 >= world: one unit per PRN
    mine = [u for u in range(len(units)) if u % world == rank]
    acqs, acq_unit = [], []
    for u in mine:
        sysc, prn = prn_list[units[u].prn_slot]
        if sysc == "G":
            a = gb.Acquisition(svc, sampled_ms=1, ms_per_code=1, samples_per_chip=49, doppler_max=5000, doppler_step=250, fs_in=fs, samples_per_ms=50000.0,
                               samples_per_code=50000.0, threshold=0.002)
            a.set_local_code(synth.sampled_code(synth.code_table("gps_l1_ca", prn), 1.023e6, fs, n_gps))
        else:
            a = gb.Acquisition(svc, sampled_ms=4, ms_per_code=4, samples_per_chip=49, doppler_max=5000, doppler_step=125, fs_in=fs, samples_per_ms=50000.0,
                               samples_per_code=200000.0, threshold=3e-4)
            a.set_local_code(synth.sampled_code(synth.code_table("gal_e1b_sinboc", prn), 2 * 1.023e6, fs, n_e1))
        acqs.append(a)
        acq_unit.append(u)
    handles = gb.acq_handles(acqs) if acqs else None
    acq_cells = 32 * 40 * n_gps + 36 * 80 * n_e1

    phases, phase_log = {}, {}

    def mark(name, t0):
        phases[name] = phases.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
        return time.perf_counter()

    def acq_step():
        t = time.perf_counter()
        if backend_inited:
            dist.broadcast(d_block, src=0)
        torch.cuda.current_stream().synchronize()
        t = mark("acq_broadcast", t)
        svc.push_device_ptr(0, d_block.data_ptr(), n_e1 + 1000, gb.ITEM_32FC, 0)
        recs = []
        if acqs:
            res = gb.acq_run_batch(acqs, 0, 0)
            t = mark("acq_push_and_search", t)
            for u, r in zip(acq_unit, res):
                recs.append(shard.acq_record(u, units[u], r.code_phase_index, r.doppler_bin, r.doppler_hz, r.peak_mag, r.input_power))
        t = mark("acq_records", t)
        if backend_inited:
            allrecs = shard.gather_records(recs, dist, dst=0, capacity=-(-len(units) // world))
        else:
            allrecs = recs
        t = mark("acq_gather", t)
        merged = None
        if rank == 0:
            merged = shard.merge_acq_winners(allrecs)  # one winner per PRN with the reference's tie-break
        mark("acq_merge", t)
        return merged

    # ------------------------------------------------------------------ tracking: C4, 1024 channels
    fs4 = 20_000_000
    step_items = 800_000  # 40 ms
    if rank == 0:
        rng = np.random.default_rng(4)
        q = np.clip(np.rint(rng.standard_normal(2 * (step_items + 100_000)) * 16.0), -128, 127).astype(np.int8)
        d_bytes = torch.from_numpy(q.view(np.uint8).copy()).to(dev)
    else:
        d_bytes = torch.zeros(2 * (step_items + 100_000), dtype=torch.uint8, device=dev)
    n_ch = 1024
    my_ch = shard.partition(n_ch, rank, world)
    kinds = [("gps_l1_ca", None, np.array([-0.5, 0.0, 0.5], np.float32), 20000, 1023, 1575.42e6, gb.CODE_REAL),
             ("gal_e1c_sinboc", "gal_e1b_sinboc", bench.TAPS, 80000, 8184, 1575.42e6, gb.CODE_REAL),
             ("bds_b1i", None, np.array([-0.5, 0.0, 0.5], np.float32), 20000, 2046, 1561.098e6, gb.CODE_REAL),
             ("glo_l1_ca", None, np.array([-0.5, 0.0, 0.5], np.float32), 20000, 511, 1602e6, gb.CODE_CPLX)]
    chans, items = [], []
    max_epochs = 40
    for c in my_ch:
        kind, dkind, taps, epoch, tlen, carrier, ck = kinds[c % 4]
        prn = 1 + (c // 4) % 30
        m = gb.Multicorrelator(svc, len(taps), tlen, code_kind=ck, item_type=gb.ITEM_8IC)
        tab = synth.code_table(kind, 1 if kind == "glo_l1_ca" else prn)
        m.set_local_code_and_taps(tab.astype(np.complex64) if ck == gb.CODE_CPLX else tab, taps)
        if dkind:
            m.set_data_code(synth.code_table(dkind, prn), 2)
        chans.append(m)
        dop = -4500.0 + 9000.0 * (c / n_ch)
        ps = np.float32(2 * np.pi * (dop + (562500.0 * ((c // 4) % 14 - 7) if kind == "glo_l1_ca" else 0.0)) / fs4)
        cs = np.float32(tlen / epoch * (1.0 + dop / carrier))
        for e in range(step_items // epoch):
            items.append((m, 1, 1000 + 37 * (c % 64) + e * epoch, float(np.float32(np.fmod(float(ps) * epoch * e, 2 * np.pi))), float(ps), 0.0,
                          float(np.float32(0.3 + 1e-3 * e)), float(cs), 0.0, epoch))
    reqs = gb.make_reqs(items) if items else None
    per_ch_epochs = [step_items // kinds[c % 4][3] for c in my_ch]
    pack_ch = np.concatenate([np.full(ne, j) for j, ne in enumerate(per_ch_epochs)]) if my_ch else np.zeros(0, int)
    pack_ep = np.concatenate([np.arange(ne) for ne in per_ch_epochs]) if my_ch else np.zeros(0, int)
    trk_chs = n_ch // 4 * (3 * 40 * 20000 + 10 * 80000)

    def trk_step():
        t = time.perf_counter()
        if backend_inited:
            dist.broadcast(d_bytes, src=0)
        torch.cuda.current_stream().synchronize()
        t = mark("trk_broadcast", t)
        svc.push_device_ptr(1, d_bytes.data_ptr(), step_items + 100_000, gb.ITEM_8IC, 0)
        local = np.zeros((len(my_ch), max_epochs, gb.OUT_STRIDE), np.complex64)
        if reqs is not None:
            out = gb.correlate_batch(svc, reqs)
            t = mark("trk_push_and_correlate", t)
            local[pack_ch, pack_ep] = out  # request order -> [channel, epoch] (index arrays built once)
        t = mark("trk_pack", t)
        if backend_inited:
            r = shard.gather_channel_outputs(my_ch, local, n_ch, dist, dst=0)
            mark("trk_gather", t)
            return r
        return local

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        phases.clear()
        if backend_inited:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            last = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if backend_inited:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        phase_log.update({k: v / steps for k, v in phases.items()})
        return float(ms.item()) / steps, last

    sampler = bench.ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    steps, warmup = max(1, args.steps), max(3, args.warmup)
    t0 = time.perf_counter()
    acq_ms, merged = timed(acq_step, steps, warmup)
    trk_ms, full = timed(trk_step, steps, warmup)
    t1 = time.perf_counter()
    if sampler:
        sampler.stop_flag = True
        sampler.join(timeout=2)
    if rank == 0:
        found = {}
        for s in sats:
            slot = (s["prn"] - 1) if s["kind"] == "gps_l1_ca" else 32 + s["prn"] - 1
            hit = merged[merged["prn_slot"] == slot]
            n = n_gps if s["kind"] == "gps_l1_ca" else n_e1
            ok = len(hit) == 1 and abs(int(hit[0]["code_phase_index"]) - s["delay_samples"] % n) <= 1 and abs(int(hit[0]["doppler_hz"]) - s["doppler_hz"]) <= 250
            found["%s%d" % ("G" if s["kind"] == "gps_l1_ca" else "E", s["prn"])] = bool(ok)
        assert len(merged) == len(prn_list), "a PRN's record is missing from the gather"
        assert full.shape == (n_ch, max_epochs, gb.OUT_STRIDE) and np.isfinite(full).all() and np.abs(full[:, 0, 1]).min() > 0
        line = {
            "metric": "acq grid cells/s (tracking channel*Msamples/s in 'tracking_value'), ONE fixed job sharded over the GPUs",
            "value": acq_cells / (acq_ms * 1e-3), "unit": "grid cells/s", "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": acq_ms + trk_ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "mode": "shard",
            "config": {"workload": "C5 cold start: 32 GPS (N=50000, 40 bins) + 36 Galileo E1 (N=200000, 80 bins) PRNs at 50 Msps gr_complex, 640 M cells per step; "
                                   "C4 tracking: 1024 channels (GPS / E1 pilot / B1I / GLONASS complex-code) on 20 Msps int8, 40 ms per step",
                       "in_timed_region": "NCCL broadcast of the sample block, device push, own share through the C ABI, NCCL all-gather of records / outputs, merge on rank 0",
                       "acq_ms_per_step": acq_ms, "trk_ms_per_step": trk_ms, "units_per_rank": len(mine), "channels_per_rank": len(my_ch)},
            "tracking_value": trk_chs / (trk_ms * 1e-3) / 1e6, "tracking_unit": "channel*Msamples/s",
            "host_phases_ms_per_step_rank0": phase_log,
            "acq_found": found, "clocks": sampler.summary(t0, t1),
        }
        print(json.dumps(line))
    if backend_inited:
        dist.destroy_process_group()
