"""Deterministic synthetic GNSS IQ and code-table access for tests and bench.

x[n] = sum_s A_s * c_s((t - tau_s) * Rc * (1 + f_s / f_carrier)) * exp(j (2 pi f_s t + phi_s)) + w[n]
with w complex white Gaussian (sigma^2 = 1 per component) from numpy PCG64(seed) and
A_s = sqrt(2 * 10^(CN0/10) / fs).  Code tables come from tests/golden/codes.npz, which was produced by
the reference's own generators (tests/golden/make_golden.py).
"""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_codes = None

CODE_SHAPES = {
    "gps_l1_ca": (32, 1023), "gal_e1b_sinboc": (36, 8184), "gal_e1c_sinboc": (36, 8184),
    "gps_l5i": (32, 10230), "gps_l5q": (32, 10230), "bds_b1i": (33, 2046), "glo_l1_ca": (1, 511),
}
# chips per second of the TABLE (E1 tables hold 2 samples per chip: sinBOC(1,1))
TABLE_RATE = {
    "gps_l1_ca": 1.023e6, "gal_e1b_sinboc": 2 * 1.023e6, "gal_e1c_sinboc": 2 * 1.023e6,
    "gps_l5i": 10.23e6, "gps_l5q": 10.23e6, "bds_b1i": 2.046e6, "glo_l1_ca": 0.511e6,
}
CARRIER_HZ = {
    "gps_l1_ca": 1575.42e6, "gal_e1b_sinboc": 1575.42e6, "gal_e1c_sinboc": 1575.42e6,
    "gps_l5i": 1176.45e6, "gps_l5q": 1176.45e6, "bds_b1i": 1561.098e6, "glo_l1_ca": 1602e6,
}


def code_table(kind, prn):
    """float32 +-1 table exactly as the reference generator returns it (1 sample/chip; 2/chip for E1 sinBOC)."""
    global _codes
    if _codes is None:
        _codes = np.load(os.path.join(GOLDEN, "codes.npz"))
    rows, length = CODE_SHAPES[kind]
    bits = np.unpackbits(_codes[kind])[: rows * length].reshape(rows, length)
    return (bits[prn - 1].astype(np.float32) * 2.0 - 1.0)


def golden_array(name):
    global _codes
    if _codes is None:
        _codes = np.load(os.path.join(GOLDEN, "codes.npz"))
    return _codes[name]


def sampled_code(table, table_rate, fs, n):
    """set_local_code-style nearest-neighbour sampling used for the local acquisition replica:
    sample i takes table[floor(i * table_rate / fs) mod L] (double arithmetic; test-side helper)."""
    idx = np.floor(np.arange(n, dtype=np.float64) * (table_rate / fs)).astype(np.int64) % len(table)
    return table[idx].astype(np.complex64)


def make_iq(fs, n, sats, seed, noise=True):
    """sats: list of dict(kind, prn, cn0_dbhz, doppler_hz, delay_samples, phase=0.0[, pilot_kind][, symbols][, symbol_periods]
    [, secondary][, doppler_rate_hz_s][, carrier_offset_hz]).
    symbols: +-1 array multiplying the data component, one entry per `symbol_periods` code periods (cyclic);
    secondary: '0'/'1' string multiplying the pilot (or, without pilot_kind, the only) component per code period, '0' -> +1;
    doppler_rate_hz_s: linear Doppler drift (carrier phase gets the quadratic term; the code rate follows the mean Doppler).
    Returns complex64[n]."""
    rng = np.random.default_rng(seed)
    if noise:
        x = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    else:
        x = np.zeros(n, np.complex128)
    t = np.arange(n, dtype=np.float64) / fs
    for s in sats:
        # "table" (+ "table_rate", "carrier_hz"): a caller-supplied +-1 code table instead of a named one (signals without a golden table)
        tab = np.asarray(s["table"], np.float64) if "table" in s else code_table(s["kind"], s["prn"]).astype(np.float64)
        rate = (s["table_rate"] if "table" in s else TABLE_RATE[s["kind"]]) * (1.0 + s["doppler_hz"] / (s["carrier_hz"] if "table" in s else CARRIER_HZ[s["kind"]]))
        tau = s["delay_samples"] / fs
        pos = np.floor((t - tau) * rate).astype(np.int64)
        idx = pos % len(tab)
        amp = np.sqrt(2.0 * 10.0 ** (s["cn0_dbhz"] / 10.0) / fs)
        sig = tab[idx]
        period = pos // len(tab)
        if "symbols" in s:
            sym = np.asarray(s["symbols"], np.float64)
            sig = sig * sym[(period // int(s.get("symbol_periods", 1))) % len(sym)]
        sec = None
        if "secondary" in s:
            sec = np.array([1.0 if ch == "0" else -1.0 for ch in s["secondary"]])[period % len(s["secondary"])]
        if "pilot_kind" in s:  # data + pilot on the same carrier (E1B + E1C: e1b - e1c, both BPSK-on-real here)
            ptab = code_table(s["pilot_kind"], s["prn"]).astype(np.float64)
            pil = ptab[idx] if sec is None else ptab[idx] * sec
            sig = (sig - pil) / np.sqrt(2.0)
        elif sec is not None:
            sig = sig * sec
        # carrier_offset_hz: a frequency offset that is NOT Doppler (GLONASS FDMA channel k * 562.5 kHz): carrier only, code rate untouched
        ph = 2 * np.pi * (s["doppler_hz"] + s.get("carrier_offset_hz", 0.0)) * t + s.get("phase", 0.0)
        if "doppler_rate_hz_s" in s:
            ph = ph + np.pi * s["doppler_rate_hz_s"] * t * t
        x += amp * sig * np.exp(1j * ph)
    return x.astype(np.complex64)


def quantize(x, item):
    """gr_complex -> interleaved int16 (x * 2^10) or int8 (x * 2^4), saturated."""
    if item == "16ic":
        q = np.clip(np.rint(np.stack([x.real, x.imag], -1) * 1024.0), -32768, 32767).astype(np.int16)
    elif item == "8ic":
        q = np.clip(np.rint(np.stack([x.real, x.imag], -1) * 16.0), -128, 127).astype(np.int8)
    else:
        raise ValueError(item)
    return np.ascontiguousarray(q.reshape(-1))


def trk16_cases():
    """Deterministic case list shared by the generator and the tests (inputs are re-created from the seed, not stored)."""
    rng = np.random.default_rng(20260016)
    cases = []
    for k in range(16):
        amp = [3.0, 40.0, 400.0, -1.0][k % 4]
        n = int(rng.integers(200, 12000))
        code_len = int(rng.choice([511, 1023, 2046]))
        n_taps = int(rng.choice([1, 3, 5]))
        if amp < 0:
            x = np.clip(np.round(rng.standard_normal(2 * n) * 5.0 + 60.0), -32768, 32767).astype(np.int16)
        else:
            x = np.clip(np.round(rng.standard_normal(2 * n) * amp), -32768, 32767).astype(np.int16)
        code = rng.choice([-1, 1], 2 * code_len).astype(np.int16)
        code[1::2] = rng.choice([-1, 0, 1], code_len)
        shifts = np.sort(rng.uniform(-1.5, 1.5, n_taps)).astype(np.float32)
        args = [float(rng.uniform(0, 6.2)), float(rng.uniform(-0.03, 0.03)), float(rng.uniform(0, 1)), float(code_len / 4000.0 * rng.uniform(0.9, 1.1))]
        if amp < 0:
            code[0::2] = 1
            code[1::2] = 0
            args[1] = 0.0
        cases.append(dict(x=x, code=code, shifts=shifts, args=[np.float32(a).item() for a in args]))
    return cases


def iq_pilot_scene(fs, n, prn, cn0_dbhz, doppler_hz, delay_samples, seed, data_flip_at=None, pilot_flip_at=None, phase=0.3):
    """A data component on I and a pilot in quadrature on one carrier, x = A (c_I d + j c_Q s) e^{j(2 pi f t + phi)} / sqrt(2) + w,
    the shape galileo_e5a_noncoherentIQ_acquisition_caf_cc searches.  The 10230-chip GPS L5 I5/Q5 tables stand in for the
    E5a-I/E5a-Q primaries (same length and rate; the block itself is code-agnostic).  `data_flip_at` / `pilot_flip_at`:
    sample index from which the component's sign is inverted (a data-bit / secondary-code transition).  complex64[n]."""
    rng = np.random.default_rng(seed)
    w = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    comp = []
    for kind, flip_at in (("gps_l5i", data_flip_at), ("gps_l5q", pilot_flip_at)):
        c = make_iq(fs, n, [dict(kind=kind, prn=prn, cn0_dbhz=cn0_dbhz - 3.0, doppler_hz=doppler_hz, delay_samples=delay_samples, phase=phase)],
                    seed=0, noise=False).astype(np.complex128)
        if flip_at is not None:
            c[int(flip_at):] *= -1.0
        comp.append(c)
    return (w + comp[0] + 1j * comp[1]).astype(np.complex64)


def iq_pilot_codes(fs, n, prn, samples_per_code=None, zero_pad_from=None):
    """Local replicas (codeI, codeQ) of iq_pilot_scene, `n` samples; optionally zeroed from `zero_pad_from` on (Zero_padding)."""
    ci = sampled_code(code_table("gps_l5i", prn), TABLE_RATE["gps_l5i"], fs, n)
    cq = sampled_code(code_table("gps_l5q", prn), TABLE_RATE["gps_l5q"], fs, n)
    if zero_pad_from is not None:
        ci[zero_pad_from:] = 0
        cq[zero_pad_from:] = 0
    return ci, cq
