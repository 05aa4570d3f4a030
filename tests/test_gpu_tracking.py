"""GPU parity: batched multicorrelator (libgnss_b200.so via the C ABI) against the oracle.

Parity metric: per request max_v |gpu_v - oracle_v| <= 1e-4 * max_v |oracle_v|, oracle =
the reference's _generic protokernels (oracle/gnss_oracle.c is bit-identical to them, see
test_oracle_cpu.py).  Chip indices must be identical, which the float64 arbiter on the oracle's own
indices (go_multicorrelator_real_f64) checks far below the 1e-4 budget.
"""
import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu
TOL = 1e-4


def rel_err(got, want):
    return float(np.abs(got - want).max() / np.abs(want).max())


def test_golden_cases(service, golden_dir):
    """Known answers produced by the reference's own Cpu_Multicorrelator_Real_Codes (tests/golden/trk_golden.npz)."""
    import gnss_b200 as gb

    g = np.load(golden_dir + "/trk_golden.npz")
    codes = [g["gps_code"], g["e1_code"]]
    for i in range(int(g["n_cases"])):
        c = {k: g["c%d_%s" % (i, k)] for k in ("code_id", "shifts", "x", "hd", "out_generic", "out_avx", "rem_carr", "phase_step",
                                                "rem_code", "code_step", "phase_rate", "code_rate")}
        code = codes[int(c["code_id"])]
        x = c["x"]
        service.push(0, x, gb.ITEM_32FC, 0)
        mc = gb.Multicorrelator(service, len(c["shifts"]), len(code), high_dyn=bool(c["hd"]))
        mc.set_local_code_and_taps(code, c["shifts"])
        got = mc.correlate(0, 0, float(c["rem_carr"]), float(c["phase_step"]), float(c["rem_code"]), float(c["code_step"]), len(x),
                           float(c["phase_rate"]), float(c["code_rate"]))
        mc.close()
        e = rel_err(got, c["out_generic"])
        assert e < TOL, (i, e, got, c["out_generic"])
        if not bool(c["hd"]):
            f64 = np.zeros(2 * len(c["shifts"]))
            ol.oracle().go_multicorrelator_real_f64(np.ascontiguousarray(x).view(np.float32), len(x), code, len(code), c["shifts"], len(c["shifts"]),
                                                    float(c["rem_carr"]), float(c["phase_step"]), float(c["rem_code"]), float(c["code_step"]), f64)
            f64 = f64.view(np.complex128)
            # identical chip indices (one flipped chip moves a tap by ~2/N = 5e-4 of the prompt); the residual is the NCO
            # frequency offset of the reference's float32-rounded phase increment, which the engine reproduces on purpose
            assert np.abs(got - f64).max() / np.abs(f64).max() < 5e-5, (i, got, f64)


@pytest.mark.parametrize("item", ["32fc", "16ic", "8ic"])
def test_batch_random_vs_oracle(service, item):
    """Many (channel, epoch) requests in one launch: ragged lengths, 1/3/5 taps, data tap, all item types."""
    import gnss_b200 as gb

    rng = np.random.default_rng(11)
    n_total = 120000
    x = (rng.standard_normal(n_total) + 1j * rng.standard_normal(n_total)).astype(np.complex64) * 3.0
    it = {"32fc": gb.ITEM_32FC, "16ic": gb.ITEM_16IC, "8ic": gb.ITEM_8IC}[item]
    if item == "32fc":
        service.push(0, x, it, 1000)
        xf = x
    else:
        q = synth.quantize(x, item)
        service.push(0, q, it, 1000)
        xf = (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)
    gps = synth.code_table("gps_l1_ca", 3)
    e1c = synth.code_table("gal_e1c_sinboc", 4)
    e1b = synth.code_table("gal_e1b_sinboc", 4)
    chans = []
    m = gb.Multicorrelator(service, 3, 1023, item_type=it)
    m.set_local_code_and_taps(gps, [-0.5, 0.0, 0.5])
    chans.append((m, gps, None, [-0.5, 0.0, 0.5], 1023 / 4000))
    m = gb.Multicorrelator(service, 5, 8184, item_type=it)
    m.set_local_code_and_taps(e1c, [-1.2, -0.3, 0.0, 0.3, 1.2])
    m.set_data_code(e1b, 2)
    chans.append((m, e1c, e1b, [-1.2, -0.3, 0.0, 0.3, 1.2], 8184 / 32000))
    m = gb.Multicorrelator(service, 1, 1023, item_type=it)
    m.set_local_code_and_taps(gps, [0.0])
    chans.append((m, gps, None, [0.0], 0.4))
    m = gb.Multicorrelator(service, 2, 1023, item_type=it)  # generic-tap-count kernel
    m.set_local_code_and_taps(gps, [-0.25, 0.25])
    chans.append((m, gps, None, [-0.25, 0.25], 0.7))
    items, meta = [], []
    for k in range(60):
        ci = k % len(chans)
        m, code, dcode, sh, step = chans[ci]
        n = int(rng.choice([1, 31, 255, 256, 257, 1000, 4000, 4096, 4097, 9000, 32000, 40001]))
        first = 1000 + int(rng.integers(0, n_total - n))
        p = (float(rng.uniform(-3.1, 3.1)), float(rng.uniform(-0.2, 0.2)), float(rng.uniform(-2.0, 2.0)), float(step * (1 + rng.uniform(-1e-5, 1e-5))))
        items.append((m, 0, first, p[0], p[1], 0.0, p[2], p[3], 0.0, n))
        meta.append((ci, first, n, p))
    out = gb.correlate_batch(service, gb.make_reqs(items))
    worst = 0.0
    for k, (ci, first, n, p) in enumerate(meta):
        m, code, dcode, sh, step = chans[ci]
        seg = xf[first - 1000: first - 1000 + n]
        want = ol.multicorrelator_real("oracle", seg, code, np.asarray(sh, np.float32), p[0], p[1], p[2], p[3])
        got = out[k, : len(sh)]
        scale = np.abs(want).max()
        if dcode is not None:
            wd = ol.multicorrelator_real("oracle", seg, dcode, np.asarray([sh[2]], np.float32), p[0], p[1], p[2], p[3])
            want = np.concatenate([want, wd])
            got = out[k, : len(sh) + 1]
            scale = max(scale, np.abs(wd).max())
        e = float(np.abs(got - want).max() / scale)
        worst = max(worst, e)
        assert e < TOL, (k, ci, n, e, got, want)
    for c in chans:
        c[0].close()
    print("worst relative error", worst)


def test_complex_codes_and_general_path(service):
    """GLONASS-style complex code tables (Cpu_Multicorrelator) and the no-window fall-back (step > 1 table sample/sample)."""
    import gnss_b200 as gb

    rng = np.random.default_rng(5)
    n = 6000
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    service.push(1, x, gb.ITEM_32FC, 0)
    code = (rng.choice([-1.0, 1.0], 511) + 1j * rng.choice([-1.0, 1.0], 511)).astype(np.complex64)
    sh = np.array([-0.5, 0.0, 0.5], np.float32)
    m = gb.Multicorrelator(service, 3, 511, code_kind=gb.CODE_CPLX)
    m.set_local_code_and_taps(code, sh)
    got = m.correlate(1, 0, 0.4, -0.03, 0.6, 511 / 6000.0, n)
    want = np.zeros(6, np.float32)
    ol.oracle().go_multicorrelator_cplx(x.view(np.float32), n, code.view(np.float32), 511, sh, 3, 0.4, -0.03, 0.6, np.float32(511 / 6000.0), want)
    assert rel_err(got, want.view(np.complex64)) < TOL
    m.close()
    gps = synth.code_table("gps_l1_ca", 9)
    m = gb.Multicorrelator(service, 3, 1023)
    m.set_local_code_and_taps(gps, sh)
    for step, rem in ((2.5, 0.3), (0.256, -5000.7), (0.256, 123456.5)):  # window too wide / far-away remnants
        got = m.correlate(1, 0, 0.1, 0.02, rem, step, n)
        want = ol.multicorrelator_real("oracle", x, gps, sh, 0.1, 0.02, rem, step)
        assert rel_err(got, want) < TOL, (step, rem)
    m.close()


def test_ring_wraparound_and_errors(service):
    """Windows that straddle the ring's wrap point are contiguous through the mirror; stale or future samples are refused."""
    import gnss_b200 as gb

    s = gb.Service(device=0, ring_capacity=10000, ring_slack=5000)
    try:
        rng = np.random.default_rng(2)
        x = (rng.standard_normal(26000) + 1j * rng.standard_normal(26000)).astype(np.complex64)
        gps = synth.code_table("gps_l1_ca", 1)
        sh = np.array([-0.5, 0.0, 0.5], np.float32)
        m = gb.Multicorrelator(s, 3, 1023)
        m.set_local_code_and_taps(gps, sh)
        for blk in range(0, 26000, 2000):
            s.push(0, x[blk: blk + 2000], gb.ITEM_32FC, blk)
        first = 18500  # positions 8500..12500 -> wraps at 10000
        got = m.correlate(0, first, 0.0, 0.01, 0.0, 0.25575, 4000)
        want = ol.multicorrelator_real("oracle", x[first: first + 4000], gps, sh, 0.0, 0.01, 0.0, 0.25575)
        assert rel_err(got, want) < TOL
        with pytest.raises(gb.Gb200Error) as ei:
            m.correlate(0, 100, 0.0, 0.01, 0.0, 0.25575, 4000)  # overwritten long ago
        assert ei.value.code == -4
        with pytest.raises(gb.Gb200Error):
            m.correlate(0, 25000, 0.0, 0.01, 0.0, 0.25575, 4000)  # not pushed yet
        m.close()
    finally:
        s.close()


@pytest.mark.parametrize("item", ["32fc", "16ic", "8ic"])
def test_inverted_spectrum_ingest(item):
    """inverted_spectrum of the data-type adapters (ishort_to_complex.cc:49-59 ...): items are conjugated in their wire type as part
    of the push, wrap-around mirror included; the most negative integer keeps its two's-complement negation (-32768 / -128)."""
    import gnss_b200 as gb

    s = gb.Service(device=0, ring_capacity=10000, ring_slack=5000)
    try:
        rng = np.random.default_rng(7)
        x = (rng.standard_normal(16000) + 1j * rng.standard_normal(16000)).astype(np.complex64) * 40.0
        it = {"32fc": gb.ITEM_32FC, "16ic": gb.ITEM_16IC, "8ic": gb.ITEM_8IC}[item]
        if item == "32fc":
            host, conj = x, np.conj(x)
        else:
            dt, lo = (np.int16, -32768) if item == "16ic" else (np.int8, -128)
            q = np.clip(np.rint(np.stack([x.real, x.imag], -1)), lo, -lo - 1).astype(dt)
            q[5::97, 1] = lo  # -lo does not exist in the type: the negation wraps back to lo
            host = q.reshape(-1)
            qc = q.copy()
            qc[:, 1] = (-q[:, 1].astype(np.int32)).astype(dt)
            conj = (qc[:, 0].astype(np.float32) + 1j * qc[:, 1].astype(np.float32)).astype(np.complex64)
        s.set_inverted_spectrum(0, True)
        step = 2000 * (1 if item == "32fc" else 2)
        for k, blk in enumerate(range(0, 16000, 2000)):
            s.push(0, host[k * step: (k + 1) * step], it, blk)
        gps = synth.code_table("gps_l1_ca", 3)
        sh = np.array([-0.5, 0.0, 0.5], np.float32)
        m = gb.Multicorrelator(s, 3, 1023, item_type=it)
        m.set_local_code_and_taps(gps, sh)
        first = 8500  # wraps at 10000: the second half is read through the mirror
        got = m.correlate(0, first, 0.3, -0.02, 0.1, 0.25575, 4000)
        want = ol.multicorrelator_real("oracle", conj[first: first + 4000], gps, sh, 0.3, -0.02, 0.1, 0.25575)
        assert rel_err(got, want) < TOL
        s.set_inverted_spectrum(0, False)
        s.push(0, host[: step], it, 16000)
        m.close()
    finally:
        s.close()


def test_int16_correlator_bit_exact(service):
    """Cpu_Multicorrelator_16sc semantics (rintf to int16, 16-bit products, per-sample saturating accumulation) reproduced
    BIT-exactly by the parallel clamp-add formulation, including runs that saturate, through single and batched calls."""
    import ctypes as C

    import gnss_b200 as gb

    i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
    f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    L = ol.oracle()
    L.go_multicorrelator_16sc.argtypes = [i16p, C.c_int, i16p, C.c_int, f32p, C.c_int] + [C.c_float] * 4 + [i16p]
    rng = np.random.default_rng(16)
    n_total = 60000
    saturated = 0
    for amp in (3.0, 40.0, 400.0, -1.0):
        if amp < 0:  # a DC stream: with an all-ones code and a frozen NCO the accumulators march into the rails
            x = np.clip(np.round(rng.standard_normal(2 * n_total) * 5.0 + 60.0), -32768, 32767).astype(np.int16)
        else:
            x = np.clip(np.round(rng.standard_normal(2 * n_total) * amp), -32768, 32767).astype(np.int16)
        first = int(rng.integers(0, 1 << 20))
        service.push(3, x, gb.ITEM_16IC, first)
        reqs_items, wants, mcs = [], [], []
        for trial in range(12):
            n_taps = int(rng.choice([1, 3, 5, 7]))
            code_len = int(rng.choice([511, 1023, 2046]))
            n = int(rng.integers(100, 20000))
            off = int(rng.integers(0, n_total - n))
            code = rng.choice([-1, 1], 2 * code_len).astype(np.int16)
            code[1::2] = rng.choice([-1, 0, 1], code_len)
            shifts = np.sort(rng.uniform(-1.5, 1.5, n_taps)).astype(np.float32)
            args = (float(rng.uniform(0, 6.2)), float(rng.uniform(-0.03, 0.03)), float(rng.uniform(0, 1)), float(code_len / 4000.0 * rng.uniform(0.9, 1.1)))
            if amp < 0:
                code[0::2] = 1 if trial % 2 else -1
                code[1::2] = 0
                args = (float(rng.uniform(0, 6.2)), 0.0 if trial % 4 else 1e-5, args[2], args[3])
            want = np.zeros(2 * n_taps, np.int16)
            L.go_multicorrelator_16sc(np.ascontiguousarray(x[2 * off: 2 * (off + n)]), n, code, code_len, shifts, n_taps, *args, want)
            mc = gb.Multicorrelator(service, n_taps, 2046, code_kind=gb.CODE_16SC, item_type=gb.ITEM_16IC)
            mc.set_local_code_and_taps(code, shifts)
            got = mc.correlate_16sc(3, first + off, args[0], args[1], args[2], args[3], n)
            assert np.array_equal(got.reshape(-1), want), (amp, trial, n, got.reshape(-1), want)
            saturated += int(np.any(np.abs(want.astype(np.int32)) >= 32767))
            reqs_items.append((mc, 3, first + off, args[0], args[1], 0.0, args[2], args[3], 0.0, n))
            wants.append(want)
            mcs.append(mc)
        outs = gb.correlate_16sc_batch(service, gb.make_reqs(reqs_items))
        for k, want in enumerate(wants):
            assert np.array_equal(outs[k].reshape(-1)[: want.size], want), (amp, k)
        # the float entry points refuse an int16 handle instead of misreading its code table
        with pytest.raises(gb.Gb200Error):
            mcs[0].correlate(3, first, 0.0, 0.0, 0.0, 0.25, 1000)
        for mc in mcs:
            mc.close()
    assert saturated >= 6


def test_service_coalesces_concurrent_channel_threads():
    """One thread per channel (the GNU Radio model) calling the blocking gb200_trk_correlate: with batch_window_us > 0 the service
    serves them with far fewer launches than calls, and every caller gets exactly the result of a lone call."""
    import threading

    import gnss_b200 as gb

    n_ch, epochs, n = 16, 40, 4000
    rng = np.random.default_rng(8)
    x = (rng.standard_normal(n * (epochs + 2)) + 1j * rng.standard_normal(n * (epochs + 2))).astype(np.complex64)
    params = [(0.1 * c, 0.002 * (c + 1), 0.05 * c, 1023.0 / 4000.0 * (1 + 1e-6 * c)) for c in range(n_ch)]
    shifts = np.array([-0.5, 0.0, 0.5], np.float32)

    def run(window):
        svc = gb.Service(ring_capacity=1 << 20, ring_slack=1 << 16, batch_window_us=window)
        try:
            svc.push(0, x, gb.ITEM_32FC, 0)
            chans = []
            for c in range(n_ch):
                m = gb.Multicorrelator(svc, 3, 1023)
                m.set_local_code_and_taps(synth.code_table("gps_l1_ca", 1 + c), shifts)
                chans.append(m)
            out = np.zeros((n_ch, epochs, 3), np.complex64)
            l0 = svc.launch_count()

            def worker(c):
                rc, ps, rcode, cs = params[c]
                for e in range(epochs):
                    out[c, e] = chans[c].correlate(0, e * n, rc + 0.01 * e, ps, rcode, cs, n)

            th = [threading.Thread(target=worker, args=(c,)) for c in range(n_ch)]
            for t in th:
                t.start()
            for t in th:
                t.join()
            launches = svc.launch_count() - l0
            stats = svc.coalescing_stats()
            for m in chans:
                m.close()
            return out, launches, stats
        finally:
            svc.close()

    alone, launches_alone, _ = run(0)
    # every service call = one request-record copy kernel + one correlator kernel
    assert launches_alone == 2 * n_ch * epochs
    for attempt in range(3):  # how many callers meet in one window depends on host thread scheduling: allow a re-run on a loaded box
        together, launches_together, (batches, requests) = run(2000)
        assert requests == n_ch * epochs and 2 * batches == launches_together
        if launches_together <= launches_alone // 3:
            break
    assert launches_together <= launches_alone // 3, (launches_together, launches_alone)
    assert np.array_equal(alone, together)


def test_tap_count_change_drops_a_stale_data_tap(service):
    """set_local_code_and_taps with fewer taps than the attached data code's prompt tap: the data tap is dropped (its prompt tap no
    longer exists) instead of reading a tap that is not there; re-attaching with a valid prompt tap brings it back."""
    import gnss_b200 as gb

    rng = np.random.default_rng(5)
    x = (rng.standard_normal(9000) + 1j * rng.standard_normal(9000)).astype(np.complex64)
    service.push(0, x, gb.ITEM_32FC, 0)
    e1c = synth.code_table("gal_e1c_sinboc", 7)
    e1b = synth.code_table("gal_e1b_sinboc", 7)
    m = gb.Multicorrelator(service, 5, 8184)
    m.set_local_code_and_taps(e1c, [-1.2, -0.3, 0.0, 0.3, 1.2])
    m.set_data_code(e1b, 4)
    assert m.n_out == 6
    m.set_local_code_and_taps(e1c, [-0.3, 0.0, 0.3])
    args = (0.4, 0.01, 0.3, 8184 / 32000)
    got = m.correlate(0, 100, *args, 8000)
    want = ol.multicorrelator_real("oracle", x[100:8100], e1c, np.asarray([-0.3, 0.0, 0.3], np.float32), *args)
    assert len(got) == 3 and np.abs(got - want).max() < TOL * np.abs(want).max()
    m.set_data_code(e1b, 1)
    got = m.correlate(0, 100, *args, 8000)
    wd = ol.multicorrelator_real("oracle", x[100:8100], e1b, np.asarray([0.0], np.float32), *args)
    assert len(got) == 4 and abs(got[3] - wd[0]) < TOL * max(np.abs(want).max(), abs(wd[0]))
    with pytest.raises(gb.Gb200Error):
        m.set_data_code(e1b, 3)
    m.close()
