"""Zero-forcing precoders for COLLINEAR users from the UNMODIFIED reference (build container only):
np.linalg.inv raises on the singular Gram matrix and antenna_array.py:243-251 retries every sub-carrier with np.linalg.pinv.
  A: K = 4, two co-located users                 (fused two-user path and the stage kernel)
  B: K = 6, three users, two of them co-located  (stage kernel, rank-2 Gram matrix of size 3)
Stored: user positions and the reference's precoder [K, U, N_sc] as complex64 (compared at 1e-5).

    python tests/golden/make_zf_singular_golden.py
"""
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

out = {}
for tag, k, pos in (("a", 4, [(100.0, 50.0, 1.5), (100.0, 50.0, 1.5)]),
                    ("b", 6, [(100.0, 50.0, 1.5), (100.0, 50.0, 1.5), (-40.0, 120.0, 1.5)])):
    mod, dist, tx, rx, arr = mg.build_objects(k, n_users=len(pos))
    rxs = [copy.deepcopy(rx) for _ in pos]
    for r, p in zip(rxs, pos):
        r.set_position(*p)
    hl = [mg.make_channel("los", arr, r).get_channel_mat_fd() for r in rxs]
    arr.set_precoding_matrix(channel_mat_fd=hl, zf_precoding=True)
    v = arr.get_precoding_mat()
    assert np.isfinite(v).all()
    arr.update_distortion(ibo_db=1.0, avg_sample_pow=mod.avg_sample_power)
    out["pos_" + tag] = np.array(pos)
    out["n_ant_" + tag] = k
    out["v_" + tag] = v.astype(np.complex64)
    out["sat_pow_" + tag] = arr.array_elements[0].impairment.sat_pow
    print(tag, v.shape, float(np.abs(v).max()))
np.savez_compressed(os.path.join(HERE, "zf_singular.npz"), **out)
