"""Confidence intervals for comparing two Monte-Carlo BER estimates (BASELINE.json north_star: "BER curves must fall inside
the reference's 95 % confidence intervals").

Both sides count bit errors over whole OFDM frames of B = 12 288 bits; errors inside a frame are strongly correlated (one
wrong 64-QAM symbol costs 1-6 bits, MCNC feeds wrong decisions back, every frame has its own receiver position), so a
binomial interval on the BIT count would be far too narrow.  The per-frame error count e_f is the i.i.d. unit:

    BER_hat = sum_f e_f / (N B),     var(BER_hat) = var(e_f) / (N B^2) = phi * p / (N B),    phi = var(e_f) / mean(e_f)

phi (the over-dispersion against a Poisson count) is estimated on the GPU from per-frame error counts of the same point
(`err_per_frame`); the reference ran the same process, so the same phi applies to its N_ref = bits_ref / B frames, where
bits_ref follows from its stop rule (mp_model.py:181-187: n_err_min = 1e6 errors or bits_sent_max = 1e7 bits per result slot,
main_mp_miso_mcnc_ber_vs_ebn0.py:58-59).  The statistic of a point is

    z = (p_got - p_ref) / sqrt(phi * p * (1 / bits_got + 1 / bits_ref)),      p = pooled BER.

A set of points passes when >= 85 % of them lie inside the 95 % interval (|z| <= 1.96; the expectation is 95 %, the slack
covers the sampling error of phi and of a few hundred points), >= 99 % inside |z| <= 3.29 and none beyond |z| = 5.
Points where either side expects fewer than MIN_ERRORS errors are not z-tested (no normal approximation); there the error
difference must stay under 5 sigma of the same model plus ten errors' worth of slack.
"""
import json
import os

import numpy as np

B = 12288
MIN_ERRORS = 100
LOG = {}


def ref_bits(p_ref, bits_max=1e7, err_min=1e6, bits_per_frame=B):
    """bits the reference spent on a result slot under its stop rule"""
    p_ref = np.asarray(p_ref, dtype=np.float64)
    bits = np.where(p_ref > 0, np.minimum(bits_max, err_min / np.maximum(p_ref, 1e-300)), bits_max)
    return np.ceil(bits / bits_per_frame) * bits_per_frame


def dispersion(err_per_frame):
    """phi per result slot from [n_frames, n_slots] per-frame error counts (>= 1: never narrower than Poisson)"""
    e = np.asarray(err_per_frame, dtype=np.float64)
    m = e.mean(axis=0)
    v = e.var(axis=0, ddof=1)
    return np.maximum(np.where(m > 0, v / np.maximum(m, 1e-300), 1.0), 1.0)


def z_scores(p_got, bits_got, p_ref, bits_ref, phi):
    p_got, p_ref = np.asarray(p_got, dtype=np.float64), np.asarray(p_ref, dtype=np.float64)
    bits_got, bits_ref = np.broadcast_to(bits_got, p_got.shape).astype(np.float64), np.broadcast_to(bits_ref, p_got.shape).astype(np.float64)
    pooled = (p_got * bits_got + p_ref * bits_ref) / (bits_got + bits_ref)
    sd = np.sqrt(np.asarray(phi) * np.maximum(pooled, 1e-300) * (1.0 / bits_got + 1.0 / bits_ref))
    z = (p_got - p_ref) / sd
    resolved = (np.minimum(p_got, p_ref) * np.minimum(bits_got, bits_ref)) >= MIN_ERRORS
    return z, resolved, sd


def reference_variance_factor(p_got, bits_got, p_ref, bits_ref, phi, axis):
    """How much noisier a shipped reference file is than its script's CURRENT (bits_sent_max, n_err_min) literals imply.

    Some files in figs/csv_results were written by older versions of their scripts, with budgets that
    are not recorded anywhere.  Their noise level can still be read off the file itself: along a densely sampled axis the
    second difference of D = p_ref - p_got cancels the smooth BER curve and any smooth systematic offset between the two
    estimates and keeps only point-to-point noise, E[d2^2] = sum_w w^2 (k var_ref + var_got), w = (1, -2, 1), with var_* from
    the variance model of this module.  Returns k >= 1 (ratio of sums over the resolved points).  The comparison keeps its
    power against every smooth deviation (bias, slope, a shifted waterfall), which is what a modelling error would produce."""
    p_got, p_ref = np.asarray(p_got, dtype=np.float64), np.asarray(p_ref, dtype=np.float64)
    bits_ref = np.broadcast_to(bits_ref, p_ref.shape).astype(np.float64)
    pooled = (p_got * bits_got + p_ref * bits_ref) / (bits_got + bits_ref)
    v_ref, v_got = phi * pooled / bits_ref, phi * pooled / bits_got
    ok = np.minimum(p_got, p_ref) * np.minimum(bits_got, bits_ref) >= MIN_ERRORS
    d = np.moveaxis(p_ref - p_got, axis, 0)
    vr, vg, okm = np.moveaxis(v_ref, axis, 0), np.moveaxis(v_got, axis, 0), np.moveaxis(ok, axis, 0)
    d2 = d[:-2] - 2 * d[1:-1] + d[2:]
    m = okm[:-2] & okm[1:-1] & okm[2:]
    num = np.sum((d2 ** 2)[m]) - np.sum((vg[:-2] + 4 * vg[1:-1] + vg[2:])[m])
    den = np.sum((vr[:-2] + 4 * vr[1:-1] + vr[2:])[m])
    return max(1.0, float(num / den)), int(m.sum())


def assert_within_ci(label, p_got, bits_got, p_ref, bits_ref, phi, min_points=1, ref_var_factor=1.0):
    # a reference file noisier than its script's literals imply (reference_variance_factor): fewer effective bits
    bits_ref = np.asarray(bits_ref, dtype=np.float64) / ref_var_factor
    z, ok, sd = z_scores(p_got, bits_got, p_ref, bits_ref, phi)
    zz = np.abs(z[ok])
    # unresolved points (a handful of errors on one side, no normal approximation): 5 sigma of the same variance model plus
    # ten errors' worth of slack on the side with fewer bits
    lo = ~ok
    bmin = np.minimum(np.broadcast_to(bits_got, np.shape(p_got)), np.broadcast_to(bits_ref, np.shape(p_got))).astype(np.float64)
    diff_lo = np.abs(np.asarray(p_got, dtype=np.float64) - np.asarray(p_ref, dtype=np.float64))[lo]
    bound = (5.0 * sd + 10.0 / bmin)[lo]
    rec = dict(points=int(ok.sum()), unresolved=int(lo.sum()),
               inside_95=float(np.mean(zz <= 1.96)) if zz.size else None,
               inside_999=float(np.mean(zz <= 3.29)) if zz.size else None, max_abs_z=float(zz.max()) if zz.size else None,
               mean_z=float(np.mean(z[ok])) if zz.size else None, median_phi=float(np.median(np.broadcast_to(phi, np.shape(p_got)))),
               median_rel_halfwidth_95=float(np.median(1.96 * sd[ok] / np.maximum(np.asarray(p_ref)[ok], 1e-300))) if zz.size else None,
               reference_variance_factor=float(ref_var_factor))
    LOG[label] = rec
    assert zz.size >= min_points, (label, rec)
    assert rec["inside_95"] >= 0.85 and rec["inside_999"] >= 0.99 and rec["max_abs_z"] <= 5.0, (label, rec)
    assert np.all(diff_lo <= bound), (label, "unresolved points outside the bound", diff_lo[diff_lo > bound], bound[diff_lo > bound])
    return rec


def dump_log():
    path = os.environ.get("MMIMO_CI_LOG")
    if path and LOG:
        old = {}
        if os.path.isfile(path):
            try:
                old = json.load(open(path))
            except Exception:
                old = {}
        old.update(LOG)
        with open(path, "w") as fh:
            json.dump(old, fh, indent=1)
