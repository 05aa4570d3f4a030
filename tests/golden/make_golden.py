"""Generate the committed golden fixtures from the REFERENCE ITSELF (oracle/_ref/libgnss_ref.so, i.e.
the reference's own sources compiled from /root/reference by oracle/Makefile).  Run in the build
container only:  python tests/golden/make_golden.py

Outputs (tests/golden/):
  codes.npz        PRN code tables produced by the reference's generators, bit-packed (+1 -> 1, -1 -> 0)
  trk_golden.npz   multicorrelator known answers: seeded inputs + outputs of the reference's
                   Cpu_Multicorrelator_Real_Codes / Cpu_Multicorrelator with _generic and _u_avx protokernels
  loop_golden.npz  discriminator / loop-filter / lock-detector outputs of the reference's scalar libs
  loop2_golden.npz Tracking_2nd_DLL_filter / Tracking_2nd_PLL_filter outputs
  trk16_golden.npz Cpu_Multicorrelator_16sc known answers (int16 samples and codes, saturating outputs), seeds only + outputs
  *.dat            the reference's own acquisition fixtures (src/tests/signal_samples), copied verbatim
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

R = ol.ref()
assert R is not None, "oracle/_ref not built (needs /root/reference)"


def pack(code):
    assert np.all(np.abs(code) == 1.0)
    return np.packbits(code > 0)


def make_codes():
    out = {}
    gps = np.zeros((32, 1023), np.float32)
    for prn in range(1, 33):
        R.gbref_gps_l1_ca_code_float(gps[prn - 1], prn, 0)
    out["gps_l1_ca"] = pack(gps)
    e1b = np.zeros((36, 8184), np.float32)
    e1c = np.zeros((36, 8184), np.float32)
    for prn in range(1, 37):
        R.gbref_galileo_e1_code_sinboc11_float(e1b[prn - 1], b"1B", prn)
        R.gbref_galileo_e1_code_sinboc11_float(e1c[prn - 1], b"1C", prn)
    out["gal_e1b_sinboc"] = pack(e1b)
    out["gal_e1c_sinboc"] = pack(e1c)
    l5i = np.zeros((32, 10230), np.float32)
    l5q = np.zeros((32, 10230), np.float32)
    for prn in range(1, 33):
        R.gbref_gps_l5i_code_float(l5i[prn - 1], prn)
        R.gbref_gps_l5q_code_float(l5q[prn - 1], prn)
    out["gps_l5i"] = pack(l5i)
    out["gps_l5q"] = pack(l5q)
    b1i = np.zeros((33, 2046), np.float32)  # the reference generator only fills PRN 1..33
    for prn in range(1, 34):
        R.gbref_beidou_b1i_code_float(b1i[prn - 1], prn, 0)
    out["bds_b1i"] = pack(b1i)
    glo = np.zeros(2 * 511, np.float32)
    R.gbref_glonass_l1_ca_code(glo, 0)
    glo = glo.view(np.complex64)
    assert np.all(glo.imag == 0)
    out["glo_l1_ca"] = pack(glo.real.copy())
    # sampled acquisition codes for the two acquisition fixtures (set_local_code inputs)
    c = np.zeros(2 * 4000, np.float32)
    R.gbref_gps_l1_ca_code_sampled(c, 1, 4000000, 0)
    out["gps_prn1_4msps_sampled"] = c.view(np.complex64).copy()
    c = np.zeros(2 * 16000, np.float32)
    R.gbref_galileo_e1_code_sampled(c, b"1B", 0, 1, 4000000, 0, 0)
    out["gal_e1b_prn1_4msps_sampled"] = c.view(np.complex64).copy()
    np.savez_compressed(os.path.join(HERE, "codes.npz"), **out)


def make_trk():
    rng = np.random.default_rng(20260001)
    gps = np.zeros(1023, np.float32)
    R.gbref_gps_l1_ca_code_float(gps, 7, 0)
    e1 = np.zeros(8184, np.float32)
    R.gbref_galileo_e1_code_sinboc11_float(e1, b"1C", 11)
    cases = []
    spec = [
        # (code, shifts, n, nominal step, high_dyn)
        (gps, [-0.5, 0.0, 0.5], 4000, 1023 / 4000, False),
        (gps, [-0.5, 0.0, 0.5], 2048, 1023 / 2048, False),
        (gps, [0.0], 777, 0.3, False),
        (e1, [-1.2, -0.3, 0.0, 0.3, 1.2], 16000, 8184 / 16000, False),
        (e1, [-1.2, -0.3, 0.0, 0.3, 1.2], 5000, 8184 / 16000, False),
        (gps, [-0.5, 0.0, 0.5], 4000, 1023 / 4000, True),
    ]
    for ci, (code, shifts, n, step, hd) in enumerate(spec):
        for rep in range(4):
            x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
            if rep % 2 == 0:  # add a real signal so prompt >> noise
                k = np.floor(np.arange(n) * step + 3.3).astype(int) % len(code)
                x += (4.0 * code[k] * np.exp(1j * (0.7 + 0.02 * np.arange(n)))).astype(np.complex64)
            p = dict(rem_carr=np.float32(rng.uniform(-3, 3)), phase_step=np.float32(0.02 + rng.uniform(-0.01, 0.01)),
                     rem_code=np.float32(rng.uniform(-1, 1) - 3.3 * (rep % 2 == 0)), code_step=np.float32(step * (1 + rng.uniform(-3e-6, 3e-6))),
                     phase_rate=np.float32(1e-8 if hd else 0.0), code_rate=np.float32(1e-10 if hd else 0.0))
            sh = np.asarray(shifts, np.float32)
            g = ol.multicorrelator_real("ref_generic", x, code, sh, float(p["rem_carr"]), float(p["phase_step"]), float(p["rem_code"]),
                                        float(p["code_step"]), float(p["phase_rate"]), float(p["code_rate"]), hd)
            a = ol.multicorrelator_real("ref_avx", x, code, sh, float(p["rem_carr"]), float(p["phase_step"]), float(p["rem_code"]),
                                        float(p["code_step"]), float(p["phase_rate"]), float(p["code_rate"]), hd)
            cases.append(dict(code_id=0 if code is gps else 1, shifts=sh, x=x, hd=hd, out_generic=g, out_avx=a, **p))
    out = {"gps_code": gps, "e1_code": e1, "n_cases": np.int32(len(cases))}
    for i, c in enumerate(cases):
        for k, v in c.items():
            out["c%d_%s" % (i, k)] = np.asarray(v)
    np.savez_compressed(os.path.join(HERE, "trk_golden.npz"), **out)


def make_loop():
    rng = np.random.default_rng(20260002)
    out = {}
    pts = (rng.standard_normal((64, 8)) * 100).astype(np.float32)
    out["pts"] = pts
    out["pll_two_quadrant"] = np.array([R.gbref_pll_cloop_two_quadrant_atan(p[0], p[1]) for p in pts])
    out["pll_four_quadrant"] = np.array([R.gbref_pll_four_quadrant_atan(p[0], p[1]) for p in pts])
    out["fll_four_quadrant"] = np.array([R.gbref_fll_four_quadrant_atan(p[0], p[1], p[2], p[3], 0.0, 0.001) for p in pts])
    out["dll_nc_e_minus_l"] = np.array([R.gbref_dll_nc_e_minus_l_normalized(p[0], p[1], p[2], p[3]) for p in pts])
    out["dll_nc_vemlp"] = np.array([R.gbref_dll_nc_vemlp_normalized(p[0:2].copy(), p[2:4].copy(), p[4:6].copy(), p[6:8].copy()) for p in pts])
    prompt = (rng.standard_normal(40) * 30 + 1000).astype(np.float32)
    out["prompt"] = prompt
    out["cn0"] = np.float32(R.gbref_cn0_svn_estimator(prompt.copy(), 20, 0.001))
    out["lock"] = np.float32(R.gbref_carrier_lock_detector(prompt.copy(), 20))
    # tracking_loop_filter_test.cc-style impulse responses
    imp = np.zeros(8, np.float32)
    imp[1] = 1.0
    for order in (1, 2, 3):
        for last in (0, 1):
            y = np.zeros(8, np.float32)
            R.gbref_tracking_loop_filter_run(0.001, 5.0, order, last, 0.0, imp, y, 8)
            out["tlf_o%d_l%d" % (order, last)] = y
    fd = rng.standard_normal(32).astype(np.float32)
    pd = (rng.standard_normal(32) * 0.1).astype(np.float32)
    out["fll_disc"], out["pll_disc"] = fd, pd
    for order in (2, 3):
        y = np.zeros(32, np.float32)
        R.gbref_fll_pll_filter_run(10.0, 35.0, order, 1234.5, fd, pd, 0.001, y, 32)
        out["fllpll_o%d" % order] = y
    np.savez_compressed(os.path.join(HERE, "loop_golden.npz"), **out)


def make_loop2():
    """Tracking_2nd_DLL_filter / Tracking_2nd_PLL_filter impulse + noise responses (the GLONASS / C_Aid trackers' filters)."""
    out = {}
    rng = np.random.default_rng(77)
    d = (0.05 * rng.standard_normal(64)).astype(np.float32)
    d[0] = 0.3
    out["disc"] = d
    for which, name in ((0, "dll"), (1, "pll")):
        for bw in (2.0, 50.0):
            y = np.zeros(64, np.float32)
            R.gbref_2nd_filter_run(which, bw, 0.001, d, y, 64)
            out["%s_bw%g" % (name, bw)] = y
    np.savez_compressed(os.path.join(HERE, "loop2_golden.npz"), **out)


def make_trk16():
    out = {}
    import synth

    for k, c in enumerate(synth.trk16_cases()):
        y = np.zeros(2 * len(c["shifts"]), np.int16)
        R.gbref_multicorrelator_16sc(c["x"], len(c["x"]) // 2, c["code"], len(c["code"]) // 2, c["shifts"], len(c["shifts"]), *c["args"], y)
        out["out_%d" % k] = y
    np.savez_compressed(os.path.join(HERE, "trk16_golden.npz"), **out)


if __name__ == "__main__":
    for f in ("GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat", "Galileo_E1_ID_1_Fs_4Msps_8ms.dat", "GSoC_CTTC_capture_2012_07_26_4Msps_4ms.dat"):
        src = os.path.join("/root/reference/src/tests/signal_samples", f)
        if os.path.exists(src):
            shutil.copyfile(src, os.path.join(HERE, f))
    make_codes()
    make_trk()
    make_loop()
    make_trk16()
    make_loop2()
    print("golden fixtures written to", HERE)
