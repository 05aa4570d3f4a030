"""GPU parity of the pcps_quicksync_acquisition_cc mirror (B200_Pcps_QuickSync_Acquisition on gb200_acq_create_folded, SURVEY.md
8f N3) against the oracle's restatement go_qs_* (itself held bit for bit to the reference block, tests/test_ref_blocks2_cpu.py)
and against the reference block compiled into oracle/_ref: folded index, Doppler, the disambiguated delay, well counter, sample
stamp and event EQUAL; statistic, power and normalised maximum within 1e-4 (float32 FFT association)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [HERE]
import host_lib as hl  # noqa: E402
import oracle_lib as ol  # noqa: E402
import synth  # noqa: E402

pytestmark = pytest.mark.gpu


def close(a, b, rtol=1e-4):
    return abs(a - b) <= rtol * max(abs(a), abs(b), 1e-30)


def scene(fs, spc, n_items, delay, doppler, cn0, seed, prn=7):
    sats = [dict(kind="gps_l1_ca", prn=prn, cn0_dbhz=cn0, doppler_hz=doppler, delay_samples=delay, phase=0.3)] if cn0 > 0 else []
    x = synth.make_iq(fs, n_items, sats, seed)
    code1 = synth.sampled_code(synth.code_table("gps_l1_ca", prn), 1.023e6, fs, spc).astype(np.complex64)
    return x, code1


@pytest.mark.parametrize("what,fs,folding,sampled_ms,delay,doppler,cn0,thr,max_dwells,bit_transition", [
    ("4 Msps F=2", 4_000_000, 2, 4, 2345, 1250.0, 50.0, 0.05, 1, 0),
    ("4 Msps F=4 delay in the last fold", 4_000_000, 4, 4, 3890, -3750.0, 52.0, 0.05, 1, 0),
    ("8 Msps F=2 general-radix FFT", 8_000_000, 2, 2, 7001, 500.0, 50.0, 0.05, 1, 0),
    ("noise only, two dwells", 4_000_000, 2, 2, 0, 0.0, -1.0, 1e9, 2, 0),
    ("bit transition rule, decision after the second dwell", 4_000_000, 2, 2, 123, -250.0, 50.0, 0.05, 2, 1),
])
def test_quicksync_block_matches_oracle_and_reference_block(service, what, fs, folding, sampled_ms, delay, doppler, cn0, thr, max_dwells, bit_transition):
    import gnss_b200 as gb
    from test_ref_blocks_cpu import quiet

    spm = fs // 1000
    spc = spm
    vec = sampled_ms * spm
    first = 16000
    x, code1 = scene(fs, spc, (max_dwells + 1) * vec, delay, doppler, cn0, 20260051 + folding + fs // 1_000_000)
    code = np.tile(code1, max(1, sampled_ms // folding))
    conf = dict(folding_factor=folding, sampled_ms=sampled_ms, doppler_max=5000, doppler_step=250, fs_in=fs, samples_per_ms=spm, samples_per_code=spc,
                threshold=thr, max_dwells=max_dwells, bit_transition_flag=bit_transition)
    service.push(3, x, gb.ITEM_32FC, first)
    ev, rec = hl.quicksync_attempt(service, 3, code, first, max_dwells + 1, **conf)
    o = ol.OracleQuickSync(**conf)
    o.set_local_code(code)
    o.restart()
    want_event = 2 if cn0 < 0 else 1
    assert ev == want_event, (what, ev, rec)
    assert len(rec) == (max_dwells if (cn0 < 0 or bit_transition) else 1)
    for k, row in enumerate(rec):
        ro = o.run(x[k * vec: (k + 1) * vec], first + (k + 1) * vec)
        assert int(row[6]) == ro.well_count and {0: 1, 1: 2, 2: 3}[int(row[0])] == ro.state, (what, k, row, ro.state)
        assert close(row[4], ro.input_power) and close(row[5], ro.mag) and close(row[3], ro.test_statistic), (what, k, row, ro.mag, ro.test_statistic)
        if cn0 > 0:
            assert int(row[1]) == ro.folded_index and row[2] == ro.acq_doppler_hz == doppler, (what, k, row, ro.folded_index)
            assert row[7] == ro.acq_delay_samples == delay and int(row[8]) == ro.acq_samplestamp, (what, k, row, ro.acq_delay_samples)
            assert close(row[9], max(ro.corr_output[:folding]), 1e-3)
    o.close()
    # the reference block itself, same vectors
    if ol.ref_blocks2() is not None:
        with quiet():
            b = ol.RefAcqVariantBlock(ol.REF_QUICKSYNC, sampled_ms, 5000, 250, fs, spm, spc, thr, max_dwells=max_dwells, bit_transition_flag=bit_transition,
                                      folding_factor=folding, prn=7)
            b.set_local_code(code)
            b.set_state(1)
            pos, k, bev = 0, 0, 0
            for _ in range(max_dwells + 3):
                r = b.work(x[pos:], (len(x) - pos) // vec)
                if r.state_before == 1:
                    row = rec[k]
                    assert close(row[3], r.test_statistics) and close(row[4], r.input_power) and int(row[6]) == r.count, (what, k, row, r.test_statistics)
                    if cn0 > 0:
                        assert row[7] == r.acq_delay_samples and row[2] == r.acq_doppler_hz and int(row[8]) - first == r.acq_samplestamp_samples, (what, k, row)
                    k += 1
                pos += r.consumed * vec
                if r.event:
                    bev = r.event
                    break
            b.close()
        assert bev == ev and k == len(rec)
