"""Golden vectors for the script variant of the two-user loop: a SINGLE-USER McncReceiver per user on the two-user
transmission (main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:563-572, :624-642), run on the UNMODIFIED reference with the
channels / bits of full_size_extras.npz.  Writes mu_su_loop.npz.      python tests/golden/make_mu_su_golden.py"""
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (imports the reference through oracle.refharness)

channel, corrector = mg.R["channel"], mg.R["corrector"]
g = np.load(os.path.join(HERE, "full_size_extras.npz"))
ibo, n_ant = 2.0, 8
out = {}
for tag, kw in (("mrt", dict(mr_precoding=True)), ("zf", dict(zf_precoding=True))):
    mod2, dist2, tx2, rx2, arr2 = mg.build_objects(n_ant, n_users=2)
    hl = [g["mu_h"][0], g["mu_h"][1]]
    rcv = []
    for u in range(2):
        ch = channel.MisoLosFd()
        ch.channel_mat_fd = hl[u]
        su = copy.deepcopy(arr2)
        su.update_n_users(n_users=1)
        su.set_precoding_matrix(channel_mat_fd=hl[u], **kw)                      # script :565-567
        su.update_distortion(ibo_db=ibo, avg_sample_pow=mod2.avg_sample_power)   # script :568-569
        rcv.append(corrector.McncReceiver(copy.deepcopy(su), copy.deepcopy(ch)))  # script :570-571
    arr2.set_precoding_matrix(channel_mat_fd=hl, **kw)
    arr2.update_distortion(ibo_db=ibo, avg_sample_pow=mod2.avg_sample_power)
    x = arr2.transmit(g["mu_bits"].astype(np.int64), out_domain_fd=True)
    for u in range(2):
        y = np.sum(x * hl[u], axis=0) / g["mu_agc_%s_u%d" % (tag, u)]             # noise free; script :629-636
        res = rcv[u].receive(n_iters_lst=[0, 1, 2], in_sig_fd=y)
        out["rx_eq_%s_u%d" % (tag, u)] = y
        out["det_%s_u%d" % (tag, u)] = np.stack([mg.bits2idx(b, 6) for b in res])
        out["err_%s_u%d" % (tag, u)] = np.array([int(np.sum(b != g["mu_bits"][u])) for b in res])
# single-antenna receivers on the two-user MU-MRT transmission: CncReceiver (main_multiuser_cnc_ber_vs_ebn0.py:124,600) and
# CncMuReceiver with the other user's symbols (corrector.py:288-345)
modulation = mg.R["modulation"]
mod2, dist2, tx2, rx2, arr2 = mg.build_objects(n_ant, n_users=2)
hl = [g["mu_h"][0], g["mu_h"][1]]
arr2.set_precoding_matrix(channel_mat_fd=hl, mr_precoding=True)
arr2.update_distortion(ibo_db=ibo, avg_sample_pow=mod2.avg_sample_power)
bits = g["mu_bits"].astype(np.int64)
x = arr2.transmit(bits, out_domain_fd=True)
sym = [modulation.modulate(mod2.constellation, mod2.n_bits_per_symbol, bits[u]) for u in range(2)]
for u in range(2):
    y = np.sum(x * hl[u], axis=0) / g["mu_agc_mrt_u%d" % u]
    mod1, dist1 = mg.build_objects(n_ant, ibo=ibo)[:2]
    rc = corrector.CncReceiver(copy.deepcopy(mod1), copy.deepcopy(dist1))
    res = rc.receive(n_iters_lst=[0, 1, 2], in_sig_fd=y)
    out["cnc_det_u%d" % u] = np.stack([mg.bits2idx(b, 6) for b in res])
    out["cnc_err_u%d" % u] = np.array([int(np.sum(b != g["mu_bits"][u])) for b in res])
    rm = corrector.CncMuReceiver(copy.deepcopy(mod1), copy.deepcopy(dist1))
    res = rm.receive(n_iters_lst=[0, 1, 2], in_sig_fd=y, other_usr_symbols=sym[1 - u])
    out["cncmu_det_u%d" % u] = np.stack([mg.bits2idx(b, 6) for b in res])
    out["cncmu_err_u%d" % u] = np.array([int(np.sum(b != g["mu_bits"][u])) for b in res])
np.savez_compressed(os.path.join(HERE, "mu_su_loop.npz"), **out)
print({k: v.tolist() for k, v in out.items() if k.startswith("err")})
