#!/usr/bin/env python
"""Host wall-clock breakdown of the tracking end-to-end step of bench.py (push blocks / stage / launch / fetch)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gnss-sdr-1_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench as B
import gnss_b200 as gb
import synth

host = torch.randint(-2000, 2000, (2 * B.STREAM_ITEMS,), dtype=torch.int16).pin_memory()
hp = host.data_ptr()
svc = gb.Service(device=0, ring_capacity=1 << 24, ring_slack=1 << 20)
chans = []
for c in range(B.N_CH):
    m = gb.Multicorrelator(svc, 5, 8184, item_type=gb.ITEM_16IC)
    m.set_local_code_and_taps(synth.code_table("gal_e1c_sinboc", 1 + c), B.TAPS)
    m.set_data_code(synth.code_table("gal_e1b_sinboc", 1 + c), 2)
    chans.append(m)
rl = B.trk_requests(0)
svc.push_ptr(0, hp, B.STREAM_ITEMS, gb.ITEM_16IC, 0)
reqs = gb.make_reqs((chans[c], 0, first, rc, ps, 0.0, rcode, cs, 0.0, B.EPOCH) for (c, first, rc, ps, rcode, cs) in rl)
svc.synchronize()
def t(f, n=10):
    f(); svc.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    svc.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
edges = [0] + [B.STREAM_ITEMS * (k + 1) // 8 for k in range(8)]
def push1(): svc.push_ptr(0, hp, B.STREAM_ITEMS, gb.ITEM_16IC, 0)
def push8():
    for k in range(8): svc.push_ptr(0, hp + 4 * edges[k], edges[k + 1] - edges[k], gb.ITEM_16IC, edges[k])
print("push 1 block  + sync: %.3f ms" % t(push1))
print("push 8 blocks + sync: %.3f ms" % t(push8))
print("correlate_batch (resident): %.3f ms" % t(lambda: gb.correlate_batch(svc, reqs)))
batch = gb.TrkBatch(svc, reqs)
print("staged batch launch+fetch: %.3f ms" % t(lambda: (batch.launch(), batch.fetch())))
print("staged batch launch only : %.3f ms" % t(lambda: batch.launch()))
def e2e():
    push8(); gb.correlate_batch(svc, reqs)
print("push8 + correlate_batch: %.3f ms" % t(e2e))
