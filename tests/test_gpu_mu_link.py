"""GPU parity of the fused two-user path (mmimo_mu_link_simulate -> k_prep_mu + k_link_frame<PA, MU>) through the C ABI
against the reference's golden two-user frame (tests/golden/full_size_extras.npz, BASELINE config 5) and the CPU oracle.
Tolerances as in test_gpu_link_parity.py: 1e-5 relative L2 for complex stages, decisions exact away from slicer ties."""
import numpy as np
import pytest

from conftest import load_golden, rel_l2
from oracle import link_oracle as lo
from test_mu_oracle_golden import mu_case, recover_noise

pytestmark = pytest.mark.gpu
TOL = 1e-5
MARGIN = 2e-4


@pytest.fixture(scope="module")
def engine():
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    eng = LinkEngine(0)
    yield eng
    eng.close()


def check_decisions(det_gpu, det_ref, cfg, min_exact=0.998):
    """fp32 vs complex128 feedback loop: a decision within MARGIN of a boundary may flip and feed back; the bulk agrees"""
    assert np.mean(det_gpu == det_ref) >= min_exact, np.mean(det_gpu == det_ref)


@pytest.mark.parametrize("tag", ["mrt", "zf"])
def test_mu_receiver_only_vs_reference(engine, tag):
    """precoder / PA / AGC from the injected channels + McncMuReceiver.receive on the reference's equalised signal"""
    g, cfg, hl, tx_idx, real = mu_case(tag)
    engine.configure(n_ant=8, channel="injected", pa="softlim", ibo_db=2.0, receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    rx_eq = np.stack([g["mu_rx_eq_%s_u%d" % (tag, u)][cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, 2, 28.0, precoding=tag, inject=dict(h=h, tx_idx=tx_idx[None], rx_eq_in=rx_eq),
                             dumps=("det", "v", "agc", "pa_p0"))
    assert rel_l2(np.transpose(out["v"][0], (1, 0, 2)), g["mu_v_" + tag]) < TOL
    assert abs(out["pa_p0"][0] / g["mu_sat_pow_" + tag] - 1) < 1e-5
    for u in range(2):
        assert rel_l2(out["agc"][0, u], g["mu_agc_%s_u%d" % (tag, u)][cfg.bins]) < TOL
        ref = g["mu_det_%s_u%d" % (tag, u)].astype(np.int64)
        check_decisions(out["det"][0, u].astype(np.int64), ref, cfg)
        assert np.all(np.abs(out["err"][u].astype(np.int64) - g["mu_err_%s_u%d" % (tag, u)]) <= 8)


@pytest.mark.parametrize("tag", ["mrt", "zf"])
def test_mu_full_frame_vs_reference(engine, tag):
    """whole frame: two-user array transmission, per-user channel, the reference's noise, AGC, receiver"""
    g, cfg, hl, tx_idx, real = mu_case(tag)
    noise = np.stack([n[cfg.bins] for n in recover_noise(g, cfg, hl, tx_idx, real, tag)])[None]
    engine.configure(n_ant=8, channel="injected", pa="softlim", ibo_db=2.0, receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, 2, 28.0, precoding=tag, inject=dict(h=h, tx_idx=tx_idx[None], noise=noise),
                             dumps=("det", "rx_eq"))
    for u in range(2):
        assert rel_l2(out["rx_eq"][0, u], g["mu_rx_eq_%s_u%d" % (tag, u)][cfg.bins]) < 2 * TOL
        check_decisions(out["det"][0, u].astype(np.int64), g["mu_det_%s_u%d" % (tag, u)].astype(np.int64), cfg)
        assert np.all(np.abs(out["err"][u].astype(np.int64) - g["mu_err_%s_u%d" % (tag, u)]) <= 8)


def test_mu_single_user_loop_vs_oracle(engine):
    """loop_single_user: oracle mu_frame(loop_single_user=True) is pinned bit-exact to the reference run of the script's
    receivers (tests/golden/mu_su_loop.npz); the GPU path is compared with both"""
    g, cfg, hl, tx_idx, real = mu_case("mrt")
    gs = load_golden("mu_su_loop")
    ref = lo.mu_frame(cfg, hl, tx_idx, None, 2, "mrt", loop_single_user=True, real=real)
    engine.configure(n_ant=8, channel="injected", pa="softlim", ibo_db=2.0, receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, 2, 300.0, precoding="mrt", loop_single_user=True,
                             inject=dict(h=h, tx_idx=tx_idx[None], noise=np.zeros((1, 2, cfg.n_sub_carr), np.complex64)),
                             dumps=("det", "rx_eq"))
    for u in range(2):
        assert rel_l2(out["rx_eq"][0, u], ref["rx_eq"][u]) < 2 * TOL
        check_decisions(out["det"][0, u].astype(np.int64), ref["det"][u], cfg)
        check_decisions(out["det"][0, u].astype(np.int64), gs["det_mrt_u%d" % u].astype(np.int64), cfg)
        assert np.all(np.abs(out["err"][u].astype(np.int64) - gs["err_mrt_u%d" % u]) <= 12)


def test_mu_device_rng_ber(engine):
    """64 antennas, device-generated LOS channels of two users, bits and noise: iterations reduce the BER for both users,
    ZF removes the inter-user interference that limits MU-MRT"""
    engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=0.0, receiver="mcnc")
    snr = float(lo.ebn0_to_snr(15.0, 2048, 2048, 64))
    ber = {}
    for tag in ("mrt", "zf"):
        out = engine.mu_simulate(64, 3, snr, precoding=tag, seed=11)
        ber[tag] = out["err"].astype(float) / out["bits"][:, None]
        assert np.all(ber[tag] > 0) and np.all(ber[tag] < 0.5)
    assert np.all(ber["zf"][:, 3] < ber["zf"][:, 0])
    assert np.all(ber["zf"][:, 0] <= ber["mrt"][:, 0] * 1.05)
    # results do not depend on how the realisations are batched
    a = engine.mu_simulate(8, 2, snr, precoding="zf", seed=5)
    b0 = engine.mu_simulate(4, 2, snr, precoding="zf", seed=5, frame_id0=0)
    b1 = engine.mu_simulate(4, 2, snr, precoding="zf", seed=5, frame_id0=4)
    assert np.array_equal(a["err"], b0["err"] + b1["err"])


@pytest.mark.parametrize("receiver, key", [("cnc", "cnc"), ("cnc_mu", "cncmu")])
def test_mu_cnc_receivers_vs_reference(engine, receiver, key):
    """single-antenna receivers of the multi-user scripts on the fused path: CncReceiver and CncMuReceiver
    (corrector.py:12-112, :248-345); golden from the unmodified reference (tests/golden/mu_su_loop.npz), noise free"""
    g, cfg, hl, tx_idx, real = mu_case("mrt")
    gs = load_golden("mu_su_loop")
    engine.configure(n_ant=8, channel="injected", pa="softlim", ibo_db=2.0, receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, 2, 300.0, precoding="mrt", receiver=receiver,
                             inject=dict(h=h, tx_idx=tx_idx[None], noise=np.zeros((1, 2, cfg.n_sub_carr), np.complex64)),
                             dumps=("det",))
    for u in range(2):
        check_decisions(out["det"][0, u].astype(np.int64), gs["%s_det_u%d" % (key, u)].astype(np.int64), cfg)
        assert np.all(np.abs(out["err"][u].astype(np.int64) - gs["%s_err_u%d" % (key, u)]) <= 12)
