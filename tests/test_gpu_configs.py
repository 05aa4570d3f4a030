"""GPU parity on the BASELINE.json configurations AS CONFIGURED (C2 exactly as bench.py runs it, C3, the C4 pieces, C5):
libgnss_b200.so through the C ABI against the restated pcps_acquisition oracle (oracle/gnss_oracle.c go_acq_*).

Bar: code_phase_index / doppler_bin / doppler_hz / decision EQUAL wherever the float32-DFT and float64-DFT
oracles agree on the cell (every present satellite is asserted to be such a case); statistics within 1e-4 relative.
The Acq_Conf values are the ones the reference's adapters derive for each signal (cited per test).
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import oracle_lib as ol
import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-4
POOL = max(2, min(16, os.cpu_count() or 2))


def close(a, b, rtol=RTOL):
    return abs(a - b) <= rtol * abs(b)


def tup(r):
    return (r.code_phase_index, r.doppler_bin, r.doppler_hz, r.decision)


def oracle_batch(cfg, codes, host_block, samp_count, cshort, arbiter=True, accumulator_lanes=0):
    """One oracle object per code, float32 DFT (the statement of the reference) and float64 DFT (the conditioning arbiter),
    run on a thread pool (ctypes releases the GIL; every oracle object owns its buffers)."""

    def one(code):
        out = []
        for f64 in ((0, 1) if arbiter else (0,)):
            o = ol.OracleAcq(cshort=int(cshort), fft_f64=f64, accumulator_lanes=accumulator_lanes, **cfg)
            o.set_local_code(code)
            r = o.run(host_block, samp_count)
            out.append((tup(r), r.test_statistic, r.peak_mag, r.input_power, r.acq_delay_samples, r.acq_doppler_hz, r.acq_samplestamp))
            o.close()
        return out

    with ThreadPoolExecutor(POOL) as ex:
        return list(ex.map(one, codes))


def compare(res, ora, present, what):
    """res: gb200 results; ora: oracle_batch output; present: indices that carry a satellite (must be well conditioned)."""
    n_cond = 0
    for k, (r, o) in enumerate(zip(res, ora)):
        o32 = o[0]
        conditioned = len(o) == 1 or o[0][0][:2] == o[1][0][:2]
        if k in present:
            assert conditioned, (what, k, "float32 / float64 oracles disagree on a present satellite", o)
        assert r.decision == o32[0][3], (what, k, tup(r), o32[0])
        if conditioned:
            n_cond += 1
            assert tup(r) == o32[0], (what, k, tup(r), o32[0])
            assert close(r.test_statistic, o32[1]) and close(r.peak_mag, o32[2]), (what, k, r.test_statistic, o32[1], r.peak_mag, o32[2])
            assert r.acq_delay_samples == o32[4] and r.acq_doppler_hz == o32[5], (what, k)
        if o32[3]:
            assert close(r.input_power, o32[3]), (what, k, r.input_power, o32[3])
        assert r.acq_samplestamp == o32[6]
    return n_cond


def run_batch(service, cfg, codes, host, item, first, offset, stream=2, oracle_cfg=None):
    """push `host` (complex64 or packed ints) at `first`, search every code in ONE gb200_acq_run_batch at first + offset"""
    import gnss_b200 as gb

    it = {"32fc": gb.ITEM_32FC, "16ic": gb.ITEM_16IC, "8ic": gb.ITEM_8IC}[item]
    service.push(stream, host, it, first)
    acqs = []
    for c in codes:
        a = gb.Acquisition(service, item_type=it, **cfg)
        a.set_local_code(c)
        acqs.append(a)
    res = gb.acq_run_batch(acqs, stream, first + offset)
    consumed, fft = acqs[0].consumed, acqs[0].fft_size
    for a in acqs:
        a.close()
    return res, consumed, fft


# ------------------------------------------------------------------------------------------------ C2 (the benched batch)
def test_c2_bench_batch_matches_oracle(service):
    """EXACTLY the acquisition batch bench.py times: build_stream(20260002) int16 at 8 Msps, 32 E1B handles (sampled sinBOC(1,1)
    replicas, 4 ms, N = 32000), +-5 kHz / 125 Hz = 80 bins, CFAR, threshold 2.5e-4, one gb200_acq_run_batch at item 0
    (gb_acq_fwd_col<16IC, GbFftStatic<160..>> + the 160x200 DIF inverse pair).  10 PRNs are present (1, 4, .., 28), 22 are absent:
    the absent ones must agree with the oracle too (decision, input power, and the noise arg-max wherever it is conditioned)."""
    import bench

    stream, sats = bench.build_stream(20260002, items=1_000_000)
    codes = bench.acq_codes()
    cfg = dict(sampled_ms=4, ms_per_code=4, samples_per_chip=8, doppler_max=bench.DOPPLER_MAX, doppler_step=bench.DOPPLER_STEP,
               fs_in=bench.FS, samples_per_ms=8000.0, samples_per_code=float(bench.N_FFT), threshold=bench.ACQ_THRESHOLD)
    res, consumed, fft = run_batch(service, cfg, codes, stream[: 2 * 200_000], "16ic", 0, 0)
    assert (consumed, fft) == (32000, 32000)
    ora = oracle_batch(cfg, codes, stream[: 2 * 32000], 32000, cshort=True)
    present = {s["prn"] - 1 for s in sats}
    n_cond = compare(res, ora, present, "C2")
    assert n_cond >= 30  # noise-only arg-max is conditioned in all but freak cases
    # and the satellites are where the scene put them: 4 ms of E1B alone (half the power) finds CN0 >= 38 dB-Hz at the right cell
    # and clears the threshold from 40 dB-Hz up; every absent PRN is negative
    for s in sats:
        r = res[s["prn"] - 1]
        if s["cn0_dbhz"] >= 38:
            assert abs(int(r.code_phase_index) - s["delay_samples"] % 32000) <= 1 and abs(r.doppler_hz - s["doppler_hz"]) <= 125, (s, tup(r))
        assert r.decision == (1 if s["cn0_dbhz"] >= 40 else 0), (s, tup(r), r.test_statistic)
    assert sum(r.decision for r in res) == 5


# ------------------------------------------------------------------------------------------------ C3
def test_c3_l5_e5a_batch_non_cfar(service):
    """C3: 10230-chip codes at 25 Msps (N = 25000 = 125 x 200), gr_complex, 12 handles in one batch with the first/second-peak
    statistic -- use_CFAR_algorithm defaults to false in galileo_e5a_pcps_acquisition.cc:82 (samples_per_chip = ceil(fs / 10.23e6)
    = 3, :118) -- and the GPS L5 adapter's own configuration in a second batch with its CFAR default
    (gps_l5i_pcps_acquisition.cc:76, 93, 130-135).  The GPS L5 I5 / Q5 tables stand in for the E5a primaries (same length and rate)."""
    fs = 25_000_000
    n = 25000
    prns = [1, 3, 6, 10, 24, 25]
    sats = [dict(kind="gps_l5i", prn=p, cn0_dbhz=53 - k, doppler_hz=[500, -1500, 2500, -3500, 4500, -4500][k], delay_samples=1111 + 3571 * k)
            for k, p in enumerate(prns)]
    sats += [dict(kind="gps_l5q", prn=p, cn0_dbhz=53 - k, doppler_hz=[-500, 1500, -2500, 3500, -4250, 4000][k], delay_samples=2222 + 3307 * k)
             for k, p in enumerate(prns)]
    x = synth.make_iq(fs, 2 * n, sats, 20260003)
    codes = [synth.sampled_code(synth.code_table(s["kind"], s["prn"]), 10.23e6, fs, n) for s in sats]
    cfg = dict(sampled_ms=1, ms_per_code=1, samples_per_chip=3, doppler_max=5000, doppler_step=250, fs_in=fs, samples_per_ms=25000.0,
               samples_per_code=25000.0, use_CFAR_algorithm_flag=0, threshold=2.0)
    off = 777
    res, consumed, fft = run_batch(service, cfg, codes, x, "32fc", 10_000, off)
    assert (consumed, fft) == (n, n)
    ora = oracle_batch(cfg, codes, x[off: off + n], 10_000 + off + n, cshort=False)
    assert compare(res, ora, set(range(12)), "C3 first/second peak") == 12
    for s, r in zip(sats, res):
        assert r.decision == 1 and abs(int(r.code_phase_index) - (s["delay_samples"] - off) % n) <= 1 and abs(r.doppler_hz - s["doppler_hz"]) <= 250, (s, tup(r))
    # GPS L5 adapter: CFAR statistic on the same block
    cfg["use_CFAR_algorithm_flag"] = 1
    cfg["threshold"] = 1e-3
    res, _, _ = run_batch(service, cfg, codes[:6], x, "32fc", 10_000, off)
    ora = oracle_batch(cfg, codes[:6], x[off: off + n], 10_000 + off + n, cshort=False)
    assert compare(res, ora, set(range(6)), "C3 CFAR") == 6


# ------------------------------------------------------------------------------------------------ C4 pieces
@pytest.mark.parametrize("what", ["gps_20000", "e1_80000"])
def test_c4_ibyte_acquisition(service, what):
    """C4: 20 Msps interleaved int8 input (GB200_ITEM_8IC: the reference converts cbyte to gr_complex in front of the block,
    gps_l1_ca_pcps_acquisition.cc:292-299, i.e. an exact int8 -> float conversion), GPS L1 C/A at N = 20000 and Galileo E1 at
    N = 80000 (4 ms).  The oracle gets the same items converted to float32.
    d_input_power: upstream VOLK's volk_32f_accumulator_s32f sums in float32, and on 8-bit items (|x|^2 are small integers) its
    _generic protokernel -- ONE sequential sum, which passes 2^25 here -- ends 4.2e-4 below the exact sum at N = 80000, while its
    a_avx / u_avx protokernels (eight interleaved partial sums, what VOLK dispatches on any x86 host) stay exact.  The two
    statements of the reference differ by more than the 1e-4 budget; the engine's tree sum is held to the AVX one here."""
    fs = 20_000_000
    if what == "gps_20000":
        n, kind, rate = 20000, "gps_l1_ca", 1.023e6
        cfg = dict(sampled_ms=1, ms_per_code=1, samples_per_chip=20, doppler_max=5000, doppler_step=250, fs_in=fs, samples_per_ms=20000.0,
                   samples_per_code=20000.0, threshold=0.004)
        prns, absent, cn0 = [2, 5, 13, 30], [7], 55
    else:
        n, kind, rate = 80000, "gal_e1b_sinboc", 2 * 1.023e6
        cfg = dict(sampled_ms=4, ms_per_code=4, samples_per_chip=20, doppler_max=5000, doppler_step=125, fs_in=fs, samples_per_ms=20000.0,
                   samples_per_code=80000.0, threshold=0.0015)
        prns, absent, cn0 = [11, 19], [3], 50
    sats = [dict(kind=kind, prn=p, cn0_dbhz=cn0 - k, doppler_hz=-3100 + 2050 * k, delay_samples=901 + 4999 * k) for k, p in enumerate(prns)]
    x = synth.make_iq(fs, n + 4000, sats, 20260004)
    q = synth.quantize(x, "8ic")
    codes = [synth.sampled_code(synth.code_table(kind, p), rate, fs, n) for p in prns + absent]
    off = 1500
    res, consumed, fft = run_batch(service, cfg, codes, q, "8ic", 64, off)
    assert (consumed, fft) == (n, n)
    host = q[2 * off: 2 * (off + n)].astype(np.float32).view(np.complex64)
    ora = oracle_batch(cfg, codes, host, 64 + off + n, cshort=False, accumulator_lanes=8)
    compare(res, ora, set(range(len(prns))), "C4 " + what)
    for s, r in zip(sats, res):
        assert r.decision == 1 and abs(int(r.code_phase_index) - (s["delay_samples"] - off) % n) <= 1, (s, tup(r))
    assert res[-1].decision == 0


def test_c4_beidou_b1i_ms_per_code_zero(service):
    """BeiDou B1I: the adapter never sets Acq_Conf::ms_per_code (beidou_b1i_pcps_acquisition.cc:73-106 -> 0, acq_conf.cc:38) nor
    samples_per_chip, so sampled_ms (1) != ms_per_code and pcps_acquisition doubles the FFT: 20000 samples consumed, N = 40000,
    code in the SECOND half of the padded replica (pcps_acquisition.cc:76-84, 255-269).  Only the first 20000 of the 40000 kept
    indices are distinct code phases; the arg-max index is whatever the reference's scan returns."""
    fs = 20_000_000
    n = 20000
    prns = [6, 11, 19]
    sats = [dict(kind="bds_b1i", prn=p, cn0_dbhz=56 - k, doppler_hz=-2300 + 1900 * k, delay_samples=333 + 6007 * k) for k, p in enumerate(prns)]
    x = synth.make_iq(fs, 2 * n, sats, 20260014)
    q = synth.quantize(x, "8ic")
    codes = [synth.sampled_code(synth.code_table("bds_b1i", p), 2.046e6, fs, n) for p in prns + [30]]
    cfg = dict(sampled_ms=1, ms_per_code=0, samples_per_chip=0, doppler_max=5000, doppler_step=250, fs_in=fs, samples_per_ms=20000.0,
               samples_per_code=20000.0, threshold=0.001)
    res, consumed, fft = run_batch(service, cfg, codes, q, "8ic", 0, 0)
    assert (consumed, fft) == (20000, 40000)
    host = q[: 2 * n].astype(np.float32).view(np.complex64)
    ora = oracle_batch(cfg, codes, host, n, cshort=False, accumulator_lanes=8)
    compare(res, ora, {0, 1, 2}, "B1I")
    for s, r in zip(sats, res):
        assert r.decision == 1 and abs(int(r.acq_delay_samples) - s["delay_samples"] % n) <= 1, (s, tup(r), r.acq_delay_samples)
    assert res[3].decision == 0


def test_c4_glonass_fdma_offsets_in_one_batch(service):
    """GLONASS L1 C/A: every handle searches its own FDMA channel, freq_offset = 562500 k Hz added to every bin's Doppler in the
    wipe-off (pcps_acquisition.cc:277-293, DFRQ1_GLO), k = -7 .. 6; six channels with different offsets share ONE batch (they
    cannot share forward spectra, so the service groups them by grid).  20 Msps, 511-chip code (one code for all satellites)."""
    fs = 20_000_000
    n = 20000
    ks = [-7, -3, 0, 2, 6, 5]
    sats = [dict(kind="glo_l1_ca", prn=1, cn0_dbhz=50 - j, doppler_hz=-4000 + 1600 * j, carrier_offset_hz=562500.0 * k, delay_samples=431 + 3001 * j)
            for j, k in enumerate(ks[:5])]  # channel k = 5 is empty
    x = synth.make_iq(fs, 2 * n, sats, 20260024)
    code = synth.sampled_code(synth.code_table("glo_l1_ca", 1), 0.511e6, fs, n)
    import gnss_b200 as gb

    service.push(2, x, gb.ITEM_32FC, 0)
    base = dict(sampled_ms=1, ms_per_code=1, samples_per_chip=40, doppler_max=5000, doppler_step=250, fs_in=fs, samples_per_ms=20000.0,
                samples_per_code=20000.0, threshold=0.0015)
    acqs = []
    for k in ks:
        a = gb.Acquisition(service, freq_offset_hz=562500 * k, **base)
        a.set_local_code(code)
        acqs.append(a)
    off = 4321
    res = gb.acq_run_batch(acqs, 2, off)
    for a in acqs:
        a.close()

    def one(k):
        out = []
        for f64 in (0, 1):
            o = ol.OracleAcq(freq_offset_hz=562500 * k, fft_f64=f64, **base)
            o.set_local_code(code)
            r = o.run(x[off: off + n], off + n)
            out.append((tup(r), r.test_statistic, r.peak_mag, r.input_power, r.acq_delay_samples, r.acq_doppler_hz, r.acq_samplestamp))
            o.close()
        return out

    with ThreadPoolExecutor(POOL) as ex:
        ora = list(ex.map(one, ks))
    compare(res, ora, set(range(4)), "GLONASS FDMA")
    # The reference accumulates the wipe-off phase in float32 (volk_gnsssdr_s32f_sincos_32fc: phase += step).  With |step| up to
    # 1.24 rad per sample the phase passes 2^14 rad inside the block, where its ulp is 2e-3 rad: the effective wipe-off frequency
    # is off by up to a few hundred Hz and smeared, more so for large |k|.  The engine reproduces that accumulator (GbWipeSeg), so
    # like the reference it finds k = -7 one or two bins off and loses k = 6 altogether; equality with the oracle is the bar.
    for s, r in list(zip(sats, res))[:4]:
        assert r.decision == 1 and abs(int(r.code_phase_index) - (s["delay_samples"] - off) % n) <= 1 and abs(r.doppler_hz - s["doppler_hz"]) <= 750, (s, tup(r))
    assert res[4].decision == 0 and res[5].decision == 0


# ------------------------------------------------------------------------------------------------ C5
def e1_scene_50msps(n_items, seed):
    fs = 50_000_000
    sats = [dict(kind="gal_e1b_sinboc", pilot_kind="gal_e1c_sinboc", prn=p, cn0_dbhz=50 - 2 * k, doppler_hz=-3300 + 2900 * k, delay_samples=12345 + 61111 * k)
            for k, p in enumerate([2, 12, 25])]
    return fs, sats, synth.make_iq(fs, n_items, sats, seed)


def test_c5_e1_200000_batch(service):
    """C5 cold start: Galileo E1 at 50 Msps, 4 ms -> N = 200000 = 400 x 500, 80 bins, CFAR; 3 present + 1 absent PRN in one batch."""
    n = 200000
    fs, sats, x = e1_scene_50msps(n + 1000, 20260005)
    prns = [s["prn"] for s in sats] + [31]
    codes = [synth.sampled_code(synth.code_table("gal_e1b_sinboc", p), 2 * 1.023e6, fs, n) for p in prns]
    cfg = dict(sampled_ms=4, ms_per_code=4, samples_per_chip=49, doppler_max=5000, doppler_step=125, fs_in=fs, samples_per_ms=50000.0,
               samples_per_code=200000.0, threshold=3e-4)
    res, consumed, fft = run_batch(service, cfg, codes, x, "32fc", 0, 500)
    assert (consumed, fft) == (n, n)
    ora = oracle_batch(cfg, codes, x[500: 500 + n], 500 + n, cshort=False)
    compare(res, ora, {0, 1, 2}, "C5 200000")
    for s, r in zip(sats, res):
        assert r.decision == 1 and abs(int(r.code_phase_index) - (s["delay_samples"] - 500) % n) <= 1, (s, tup(r))
    assert res[3].decision == 0


def test_c5_e1_800000_bit_transition(service):
    """The bit_transition_flag=true variant of conf/gnss-sdr_galileo_E1_extended_correlator_byte.conf:76-82 at 50 Msps:
    400000 samples consumed (2 x 4 ms), FFT 4 x 200000 = 800000 points, replica [code | 0] in the second half of the FFT input, magnitudes kept from
    outbuf[400000 .. 800000) (pcps_acquisition.cc:113-117, 730), use_CFAR_algorithm=false -> first/second-peak statistic on the
    kept rows, input power over all N incl. padding (:703-709).  One present PRN with a data-bit flip inside the block."""
    n_cons, n = 400000, 800000
    fs, sats, x = e1_scene_50msps(n_cons + 64, 20260055)
    x = x.copy()
    x[sats[0]["delay_samples"] + 200000:] *= -1.0  # a symbol transition of every component at PRN 2's code boundary
    code_1 = synth.sampled_code(synth.code_table("gal_e1b_sinboc", sats[0]["prn"]), 2 * 1.023e6, fs, 200000)
    # the adapter's replica buffer: vector_length = 2 x 200000 items of which sampled_ms / 4 = 1 period is filled, the rest stays
    # value-initialised (galileo_e1_pcps_ambiguous_acquisition.cc:123-128, 276-279)
    code = np.concatenate([code_1, np.zeros(200000, np.complex64)])
    cfg = dict(sampled_ms=4, ms_per_code=4, samples_per_chip=49, doppler_max=5000, doppler_step=125, fs_in=fs, samples_per_ms=50000.0,
               samples_per_code=200000.0, bit_transition_flag=1, use_CFAR_algorithm_flag=0, threshold=2.0)
    res, consumed, fft = run_batch(service, cfg, [code], x, "32fc", 0, 10)
    assert (consumed, fft) == (n_cons, n)
    ora = oracle_batch(cfg, [code], x[10: 10 + n_cons], 10 + n_cons, cshort=False, arbiter=False)
    compare(res, ora, {0}, "C5 800000 bit transition")
    # both code periods of the block leave a peak in the kept half, 200000 cells apart, so the first/second-peak ratio is ~1 and the
    # reference declares negative with any threshold > 1: that is the block's behaviour in this configuration, reproduced
    assert res[0].decision == 0 and int(res[0].code_phase_index) % 200000 in (12335, 12336) and res[0].doppler_hz == -3250
