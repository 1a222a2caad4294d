"""GPU parity of the fused two-user path (mmimo_mu_link_simulate -> k_prep_mu + the two-user frame kernel) through the C ABI
against the reference's golden two-user frames (tests/golden/full_size_extras.npz: K = 8; tests/golden/mu_k64.npz: K = 64,
BASELINE config 5) and the CPU oracle.  Tolerances as in test_gpu_link_parity.py: 1e-5 relative L2 for complex stages;
decisions and error counts equal to the reference's except where the oracle's slicer input lies within MARGIN of a
boundary (tests/parity_util.py) - the same rule as for the single-user link."""
import numpy as np
import pytest

from conftest import load_golden, rel_l2
from oracle import link_oracle as lo
from parity_util import check_frame
from test_mu_oracle_golden import mu_case, mu_case_k64, recover_noise

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def engine():
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    eng = LinkEngine(0)
    yield eng
    eng.close()


def check_user(case, out, u, ref, golden_det=None, golden_err=None, iters=None):
    """decisions / error counts of user u of realisation 0 against the oracle frame `ref` (lo.mu_frame, pinned bit-exactly to
    the reference in test_mu_oracle_golden.py) under the margin rule; where the frame did not diverge, also against the
    reference's stored decisions"""
    n_ok = check_frame(case, out["det"][0, u], ref["det"][u], ref["corr"][u], 64, out["err"][u], ref["err"][u])
    if golden_det is not None:
        iters = np.arange(golden_det.shape[0]) if iters is None else np.asarray(iters)
        sel = [k for k, itv in enumerate(iters) if itv < n_ok]
        assert np.array_equal(out["det"][0, u].astype(np.int64)[iters[sel]], golden_det.astype(np.int64)[sel])
        assert np.array_equal(out["err"][u].astype(np.int64)[iters[sel]], np.asarray(golden_err)[sel])
    return n_ok


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
    ref = lo.mu_frame(cfg, hl, tx_idx, None, 2, tag, real=real, rx_eq_in=rx_eq[0])
    for u in range(2):
        assert rel_l2(out["agc"][0, u], g["mu_agc_%s_u%d" % (tag, u)][cfg.bins]) < TOL
        check_user("mu K=8 %s receiver-only" % tag, out, u, ref, g["mu_det_%s_u%d" % (tag, u)], g["mu_err_%s_u%d" % (tag, u)])


@pytest.mark.parametrize("tag", ["mrt", "zf"])
def test_mu_full_frame_vs_reference(engine, tag):
    """whole frame: two-user array transmission, per-user channel, the reference's noise, AGC, receiver"""
    g, cfg, hl, tx_idx, real = mu_case(tag)
    noise = np.stack([n[cfg.bins] for n in recover_noise(g, cfg, hl, tx_idx, real, tag)])[None]
    engine.configure(n_ant=8, channel="injected", pa="softlim", ibo_db=2.0, receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, 2, 28.0, precoding=tag, inject=dict(h=h, tx_idx=tx_idx[None], noise=noise),
                             dumps=("det", "rx_eq"))
    ref = lo.mu_frame(cfg, hl, tx_idx, recover_noise(g, cfg, hl, tx_idx, real, tag), 2, tag, real=real)
    for u in range(2):
        assert rel_l2(out["rx_eq"][0, u], g["mu_rx_eq_%s_u%d" % (tag, u)][cfg.bins]) < 2 * TOL
        check_user("mu K=8 %s frame" % tag, out, u, ref, g["mu_det_%s_u%d" % (tag, u)], g["mu_err_%s_u%d" % (tag, u)])


@pytest.mark.parametrize("tag", ["mrt", "zf"])
def test_mu_single_user_loop_vs_oracle(engine, tag):
    """loop_single_user: oracle mu_frame(loop_single_user=True) is pinned bit-exact to the reference run of the script's
    receivers (tests/golden/mu_su_loop.npz), MU-MRT and - with the phase-only PA power the script ends up with - ZF; the GPU
    path is compared with both"""
    g, cfg, hl, tx_idx, real = mu_case(tag)
    gs = load_golden("mu_su_loop")
    ref = lo.mu_frame(cfg, hl, tx_idx, None, 2, tag, loop_single_user=True, real=real)
    engine.configure(n_ant=8, channel="injected", pa="softlim", ibo_db=2.0, receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, 2, 300.0, precoding=tag, loop_single_user=True,
                             inject=dict(h=h, tx_idx=tx_idx[None], noise=np.zeros((1, 2, cfg.n_sub_carr), np.complex64)),
                             dumps=("det", "rx_eq"))
    for u in range(2):
        assert rel_l2(out["rx_eq"][0, u], ref["rx_eq"][u]) < 2 * TOL
        check_user("mu K=8 %s single-user loop" % tag, out, u, ref, gs["det_%s_u%d" % (tag, u)], gs["err_%s_u%d" % (tag, u)])


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
    ref = lo.mu_frame(cfg, hl, tx_idx, None, 2, "mrt", real=real, receiver=receiver)
    for u in range(2):
        check_user("mu K=8 mrt %s" % receiver, out, u, ref, gs["%s_det_u%d" % (key, u)], gs["%s_err_u%d" % (key, u)])


@pytest.mark.parametrize("tag", ["mrt", "zf"])
def test_mu_k64_full_frame_vs_reference(engine, tag):
    """BASELINE config 5 at its named size: 64 antennas, two users, MU-MRT and ZF, the reference's bits and noise injected,
    McncMuReceiver iterations 0..3 (tests/golden/mu_k64.npz from the unmodified reference)"""
    g, cfg, hl, tx_idx, real, noise = mu_case_k64(tag)
    n_iter = int(g["iters"].max())
    ref = lo.mu_frame(cfg, hl, tx_idx, noise, n_iter, tag, real=real)
    engine.configure(n_ant=cfg.n_ant, channel="injected", pa="softlim", ibo_db=float(g["ibo_db"]), receiver="mcnc")
    h = np.stack([hl[u][:, cfg.bins] for u in range(2)])[None]
    out = engine.mu_simulate(1, n_iter, float(g["snr_db"]), precoding=tag,
                             inject=dict(h=h, tx_idx=tx_idx[None], noise=noise[None][:, :, cfg.bins]),
                             dumps=("det", "rx_eq", "v", "agc", "pa_p0"))
    K = cfg.n_ant
    assert rel_l2(np.transpose(out["v"][0], (1, 0, 2))[[0, K - 1]], g["v_rows_" + tag]) < TOL
    assert abs(out["pa_p0"][0] / g["sat_pow_" + tag] - 1) < 1e-5
    for u in range(2):
        assert rel_l2(out["agc"][0, u], g["agc_sc_%s_u%d" % (tag, u)]) < TOL
        assert rel_l2(out["rx_eq"][0, u], g["rx_eq_sc_%s_u%d" % (tag, u)]) < 2 * TOL
        check_user("mu K=64 %s frame" % tag, out, u, ref, g["det_%s_u%d" % (tag, u)], g["err_%s_u%d" % (tag, u)], g["iters"])


def test_mu_cnc_loop_kernel_matches_single_kernel_form(monkeypatch):
    """Two-user runs with the single-antenna receivers (CncReceiver, CncMuReceiver): pass 0 in the two-user frame kernel + the
    receiver loops in k_cnc_loop_t64 (one instance per 64-thread group, the default) against whole instances in the frame
    kernel (MMIMO_CNC_SPLIT=0): identical decisions and counters of both users, odd realisation counts included."""
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    engines = {}
    try:
        for split in ("1", "0"):
            monkeypatch.setenv("MMIMO_CNC_SPLIT", split)
            engines[split] = LinkEngine(0)
        monkeypatch.delenv("MMIMO_CNC_SPLIT", raising=False)
        for receiver, tag, pa, n_ant, n_real, n_iter in (("cnc", "mrt", "softlim", 8, 7, 3), ("cnc_mu", "zf", "softlim", 16, 301, 2),
                                                         ("cnc_mu", "mrt", "rapp", 4, 5, 4), ("cnc", "zf", "toi", 64, 75, 5)):
            outs = {}
            for split, eng in engines.items():
                eng.configure(n_ant=n_ant, channel="los", pa=pa, ibo_db=1.0, receiver="mcnc")
                outs[split] = eng.mu_simulate(n_real, n_iter, 18.0, precoding=tag, receiver=receiver, seed=23, dumps=("det", "rx_eq"))
            a, b = outs["1"], outs["0"]
            for key in ("err", "det", "rx_eq"):
                assert np.array_equal(a[key], b[key]), (receiver, tag, pa, key)
            assert a["err"].sum() > 0
    finally:
        for eng in engines.values():
            eng.close()
