"""The decision-parity rule of every GPU test, in one place (BASELINE.json north_star: "hard-decision bits and error
counts must match bit-exactly"; fp32 compute against the reference's complex128):

  * per frame, iterations are compared in order; all decisions and the error count of an iteration must be EQUAL to the
    oracle's (which is pinned bit-exactly to the unmodified reference);
  * a differing decision is accepted only where the ORACLE's slicer input lies closer than MARGIN to a decision boundary
    (an fp32 result may legitimately fall on the other side); because CNC / MCNC feed decisions back, that frame is then
    compared up to that iteration only ("diverged at iteration it");
  * every comparison is tallied; with MMIMO_MARGIN_LOG=<path> the tallies are written as JSON at the end of the session
    (committed from a B200 run as profiles/r02_margin_histogram.json).
"""
import json
import os

import numpy as np

from oracle import link_oracle as lo

MARGIN = 2e-4   # absolute distance to a slicer boundary (constellation spacing 2) below which an fp32 flip is tolerated

STATS = {}


def _case(name):
    return STATS.setdefault(name, dict(frames=0, iterations_compared=0, decisions_compared=0, decisions_flipped=0,
                                       max_margin_of_a_flip=0.0, frames_diverged=0, diverged_at_iteration=[],
                                       min_margin_seen=None))


def check_frame(case, det_gpu, det_ref, corr_ref, constel, err_gpu=None, err_ref=None, margin=MARGIN):
    """det_*: [n_it, N_sc] symbol indices, corr_ref: the oracle's slicer inputs [n_it, N_sc]; err_*: [n_it] bit errors.
    Returns the number of iterations compared exactly (== n_it when the frame never diverged)."""
    st = _case(case)
    st["frames"] += 1
    det_gpu, det_ref = np.asarray(det_gpu).astype(np.int64), np.asarray(det_ref).astype(np.int64)
    n_it = det_ref.shape[0]
    for it in range(n_it):
        m_all = lo.detection_margin(corr_ref[it], constel)
        mn = float(m_all.min())
        st["min_margin_seen"] = mn if st["min_margin_seen"] is None else min(st["min_margin_seen"], mn)
        st["iterations_compared"] += 1
        st["decisions_compared"] += det_ref.shape[1]
        mism = det_gpu[it] != det_ref[it]
        if mism.any():
            m = m_all[mism]
            st["decisions_flipped"] += int(mism.sum())
            st["max_margin_of_a_flip"] = max(st["max_margin_of_a_flip"], float(m.max()))
            st["frames_diverged"] += 1
            st["diverged_at_iteration"].append(it)
            assert m.max() < margin, ("%s: iteration %d: %d decisions differ from the reference with margin up to %.3g"
                                      % (case, it, int(mism.sum()), float(m.max())))
            return it
        if err_gpu is not None:
            assert int(err_gpu[it]) == int(err_ref[it]), (case, it, int(err_gpu[it]), int(err_ref[it]))
    return n_it


def dump_stats():
    path = os.environ.get("MMIMO_MARGIN_LOG")
    if path and STATS:
        old = {}
        if os.path.isfile(path):
            try:
                old = json.load(open(path))
            except Exception:
                old = {}
        old.update(STATS)
        with open(path, "w") as fh:
            json.dump(dict(margin=MARGIN, rule=__doc__.split("\n\n")[1].strip(), cases=old) if "cases" not in old else old,
                      fh, indent=1)
