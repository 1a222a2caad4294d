"""GPU parity of the fused link kernels (through the C ABI) against the reference's own outputs
(tests/golden/*.npz) and the CPU oracle, with the reference's bits / receiver positions / channel /
noise injected.  Tolerances: complex stages 1e-5 relative L2 (fp32 compute vs complex128, BASELINE.json
north_star); hard decisions and error counts bit-exact, where a frame is only compared up to the first
iteration in which the ORACLE's slicer input sits closer than MARGIN to a decision boundary (an fp32
result may legitimately fall on the other side there, and MCNC feeds the decision back)."""
import numpy as np
import pytest

from conftest import load_golden, rel_l2
from oracle import link_oracle as lo
from parity_util import MARGIN, check_frame

pytestmark = pytest.mark.gpu

TOL = 1e-5      # relative L2, complex stages

GOLDEN_4096 = ["link_k64_los_mcnc", "link_k8_two_path_mcnc_clean", "link_k4_rayleigh_cnc", "link_k8_los_csi",
               "link_k8_los_rapp",
               # the headline shape with 32 frames and its CNC twin with 16 (tests/golden/make_golden_k64.py)
               "link_k64_los_mcnc32", "link_k64_los_cnc16",
               # BASELINE configs[2] at its named size: K=64 two-path and Rayleigh channels, 8 frames each
               "link_k64_two_path_mcnc8", "link_k64_rayleigh_mcnc8"]
# Frames that may hit a sub-margin decision, per case.  Measured on the B200 (profiles/r02_margin_histogram.json): of 1.74 M
# decisions compared over all cases NONE differed from the reference, although the oracle's slicer input came as close as
# 5e-7 to a boundary; the allowance below is slack for other inputs, the rule itself (parity_util.check_frame: a differing
# decision must sit within MARGIN of a boundary) is never relaxed.
MAX_DIVERGED_FRACTION = 1.0 / 16


@pytest.fixture(scope="module")
def engine():
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    eng = LinkEngine(0)
    yield eng
    eng.close()


def cfg_from_golden(g):
    pa = {"softlim": lo.PA_SOFT_LIMITER, "rapp": lo.PA_RAPP, "toi": lo.PA_TOI}[str(g["pa"])]
    return lo.LinkConfig(constel_size=int(g["constel"]), n_fft=int(g["n_fft"]), n_sub_carr=int(g["n_sc"]),
                         n_ant=int(g["n_ant"]), pa_kind=pa, ibo_db=float(g["ibo_db"]), channel=str(g["chan"]))


def oracle_run(cfg, snr_db, n_iter, seeds, n_frames, is_mcnc, incl_clean, csi):
    """Oracle frames incl. everything that gets injected into the GPU path."""
    # the clean loop consumes the RNG streams first (mp_model.py:136-175); the oracle handles that
    return lo.link_simulate(cfg, snr_db, n_iter, seeds, n_frames, is_mcnc=is_mcnc, incl_clean_run=incl_clean,
                            csi_eps=csi, return_frames=True)


def run_case(engine, cfg, snr_db, iters, seeds, n_frames, is_mcnc, incl_clean, csi, golden=None, name=None):
    n_iter = int(np.max(iters))
    errs, bits, frames = oracle_run(cfg, snr_db, n_iter, seeds, n_frames, is_mcnc, incl_clean, csi)
    pa = {lo.PA_SOFT_LIMITER: "softlim", lo.PA_RAPP: "rapp", lo.PA_TOI: "toi"}[cfg.pa_kind]
    engine.configure(constel_size=cfg.constel_size, n_ant=cfg.n_ant, pa=pa, ibo_db=cfg.ibo_db,
                     p_hardness=cfg.p_hardness, channel="injected", receiver="mcnc" if is_mcnc else "cnc")
    bins = cfg.bins
    inject = dict(h_true=np.stack([f["real"].h_true[:, bins] for f in frames]),
                  tx_idx=np.stack([f["tx_idx"] for f in frames]),
                  noise=np.stack([f["noise"][bins] for f in frames]))
    if csi is not None:
        inject["h_loop"] = np.stack([f["real"].h_loop[:, bins] for f in frames])
    out = engine.simulate(n_frames, n_iter, snr_db, inject=inject,
                          dumps=("det", "err_per_frame", "rx_eq", "corr", "v", "agc", "pa_p0", "alpha"))
    n_div = 0
    for i, fr in enumerate(frames):
        real = fr["real"]
        assert rel_l2(out["v"][i], real.v) < TOL
        assert rel_l2(out["agc"][i], real.ak_hk_vk_agc) < TOL
        assert abs(out["pa_p0"][i] / real.pa[1] - 1) < TOL
        assert np.allclose(out["alpha"][i], real.alpha_k, rtol=1e-5)
        assert rel_l2(out["rx_eq"][i], fr["rx_eq_sc"]) < TOL
        corr_ref = fr["corr"]
        case = name or "fused K=%d %s %s" % (cfg.n_ant, cfg.channel, "mcnc" if is_mcnc else "cnc")
        n_ok = check_frame(case, out["det"][i], fr["det"], corr_ref, cfg.constel_size, out["err_per_frame"][i], fr["err"])
        n_div += n_ok < fr["det"].shape[0]
        for it in range(n_ok):
            assert rel_l2(out["corr"][i, it], corr_ref[it]) < 5 * TOL
        if golden is not None:  # the reference's own outputs
            gdet = golden["det"][i].astype(np.int64)
            sel = [k for k, itv in enumerate(iters) if itv < n_ok]
            assert np.array_equal(out["det"][i][np.asarray(iters)[sel]], gdet[sel])
            assert rel_l2(out["rx_eq"][i], golden["rx_eq_sc"][i]) < TOL
            if i < len(golden["agc_sc"]):
                assert rel_l2(out["agc"][i], golden["agc_sc"][i]) < TOL
    if n_div == 0:
        off = 1 if incl_clean else 0
        assert np.array_equal(out["err"].astype(np.int64), errs[off:])
        if golden is not None:
            assert np.array_equal(out["err"].astype(np.int64)[np.asarray(iters)], golden["err"].astype(np.int64)[off:])
    assert n_div <= max(1, int(n_frames * MAX_DIVERGED_FRACTION)), "too many frames hit a sub-margin decision"
    return out, frames


@pytest.mark.parametrize("name", GOLDEN_4096)
def test_golden_link_cases(engine, name):
    g = load_golden(name)
    cfg = cfg_from_golden(g)
    csi = float(g["csi_eps"])
    run_case(engine, cfg, float(g["snr_db"]), [int(v) for v in g["iters"]], list(g["seeds"]), int(g["n_frames"]),
             bool(g["is_mcnc"]), bool(g["incl_clean"]), None if csi < 0 else csi, golden=g, name=name)


@pytest.mark.parametrize("constel,pa,ibo", [(4, lo.PA_SOFT_LIMITER, -3.0), (16, lo.PA_SOFT_LIMITER, 2.0),
                                            (256, lo.PA_SOFT_LIMITER, 0.0), (64, lo.PA_TOI, 20.0)])
def test_other_constellations_and_pa(engine, constel, pa, ibo):
    cfg = lo.LinkConfig(constel_size=constel, n_ant=4, pa_kind=pa, ibo_db=ibo, channel="two_path")
    snr = float(lo.ebn0_to_snr(18.0, cfg.n_sub_carr, cfg.n_sub_carr, constel))
    run_case(engine, cfg, snr, [0, 1, 2, 3], [5, 6, 7], 2, True, False, None)


def test_single_antenna(engine):
    cfg = lo.LinkConfig(n_ant=1, channel="los", ibo_db=1.0)
    snr = float(lo.ebn0_to_snr(16.0, cfg.n_sub_carr, cfg.n_sub_carr, 64))
    run_case(engine, cfg, snr, [0, 1, 2], [9, 8, 7], 2, True, False, None)
    run_case(engine, cfg, snr, [0, 1, 2], [9, 8, 7], 2, False, False, None)


def test_device_channel_generation_matches_oracle(engine):
    """LOS / two-path channel generated on the device from injected receiver positions."""
    for chan in ("los", "two_path"):
        cfg = lo.LinkConfig(n_ant=8, channel=chan)
        rng = np.random.default_rng(3)
        pos = np.stack([lo.draw_rx_position(cfg, rng) for _ in range(3)])
        engine.configure(n_ant=8, channel=chan)
        out = engine.simulate(3, 0, 20.0, inject=dict(rx_pos=pos), dumps=("h_true", "rx_pos"))
        for i in range(3):
            h = lo.draw_channel(cfg, pos[i])[:, cfg.bins]
            assert rel_l2(out["h_true"][i], h) < TOL


def test_clean_run(engine):
    g = load_golden("link_k8_two_path_mcnc_clean")
    cfg = cfg_from_golden(g)
    n = int(g["n_frames"])
    snr = float(g["snr_db"])
    # clean frames are the first n realisations of the streams
    bit_rng, noise_rng, loc_rng = (np.random.default_rng(int(s)) for s in g["seeds"])
    pos, tx, nz = [], [], []
    for _ in range(n):
        p = lo.draw_rx_position(cfg, loc_rng)
        real = lo.make_realisation(cfg, lo.draw_channel(cfg, p))
        pos.append(p)
        tx.append(lo.bits_to_indices(bit_rng.choice((0, 1), cfg.n_bits_per_ofdm_sym), 6))
        nz.append(lo.awgn_noise(noise_rng, cfg.n_fft, cfg.avg_symbol_power * real.hk_vk_noise_scaler, snr)[cfg.bins])
    engine.configure(n_ant=cfg.n_ant, channel="two_path", ibo_db=cfg.ibo_db)
    out = engine.simulate(n, 0, snr, incl_clean_run=True,
                          inject=dict(clean_rx_pos=np.array(pos), clean_tx_idx=np.array(tx), clean_noise=np.array(nz)))
    assert int(out["clean_err"][0]) == int(g["err"][0])


def test_error_messages(engine):
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import MmimoError
    with pytest.raises(MmimoError):
        engine.configure(constel_size=32)
    with pytest.raises(MmimoError):
        engine.configure(n_fft=1024, n_sub_carr=512)
    engine.configure(channel="injected")
    with pytest.raises(MmimoError):
        engine.simulate(1, 1, 10.0)


def test_fused_sdr_matches_reference_snapshots(engine):
    """mmimo_sdr_simulate vs the reference's SDR snapshot computation (tests/golden/full_size_extras.npz, generated from
    main_beampatterns_plotting/sdr_mp.py's formulae on the unmodified reference classes), receiver positions and bits injected."""
    g = load_golden("full_size_extras")
    for i, ibo in enumerate(g["sdr_ibo"]):
        engine.configure(n_ant=8, channel="los", pa="softlim", ibo_db=float(ibo))
        idx = lo.bits_to_indices(g["sdr_bits"][i], 6)
        des, dis = engine.sdr(1, inject=dict(rx_pos=g["sdr_pos"][i:i + 1], tx_idx=idx[None]))
        assert abs(des[0] / g["sdr_desired"][i] - 1) < 2e-5
        assert abs(dis[0] / g["sdr_distortion"][i] - 1) < 3e-4   # difference of nearly equal signals in fp32
    g64 = load_golden("sdr_k64")   # the size BASELINE config 4 names: 64 antennas (tests/golden/make_golden_k64.py)
    for i, ibo in enumerate(g64["sdr_ibo"]):
        engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=float(ibo))
        idx = lo.bits_to_indices(g64["sdr_bits"][i], 6)
        des, dis = engine.sdr(1, inject=dict(rx_pos=g64["sdr_pos"][i:i + 1], tx_idx=idx[None]))
        assert abs(des[0] / g64["sdr_desired"][i] - 1) < 2e-5
        assert abs(dis[0] / g64["sdr_distortion"][i] - 1) < 3e-4
    # device RNG: SDR grows with the back-off (BASELINE config 4, "SDR vs IBO")
    sdr = []
    for ibo in (0.0, 3.0, 6.0):
        engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=ibo)
        des, dis = engine.sdr(64, seed=3)
        sdr.append(10 * np.log10(np.mean(des / dis)))
    assert sdr[0] < sdr[1] < sdr[2] and 10 < sdr[0] < 25


def test_multi_point_sweep_equals_point_by_point(engine):
    """mmimo_link_sweep: one batch carrying several (Eb/N0, IBO) points gives exactly the counters of separate runs of the
    same global frame ids (the RNG is keyed on the frame id, the PA / AGC / noise scalers are per frame)"""
    engine.configure(n_ant=16, channel="los", pa="softlim", ibo_db=1.0, receiver="mcnc")
    snr = np.array([14.0, 18.0, 22.0, 18.0])
    ibo = np.array([0.0, 0.0, 3.0, 6.0])
    fpp, n_iter = 24, 3
    err = engine.sweep(snr, ibo, fpp, n_iter, seed=77, frame_id0=1000)
    for p in range(len(snr)):
        engine.set_ibo(float(ibo[p]))
        ref = engine.simulate(fpp, n_iter, float(snr[p]), seed=77, frame_id0=1000 + p * fpp)["err"]
        assert np.array_equal(err[p], ref), (p, err[p], ref)
    assert err[1, 0] < err[0, 0]          # more Eb/N0, fewer errors
    # IBO left at the configured value when no table is given
    engine.set_ibo(2.0)
    e2 = engine.sweep(snr[:2], None, fpp, n_iter, seed=5)
    r0 = engine.simulate(fpp, n_iter, float(snr[0]), seed=5, frame_id0=0)["err"]
    assert np.array_equal(e2[0], r0)
    # CncReceiver points (per-point Eb/N0 only): the receiver loop runs in k_cnc_loop_t64, which counts per point as well;
    # 23 frames per point leave groups of the last CTA round without a frame and put point boundaries inside a CTA
    engine.configure(n_ant=8, channel="los", pa="softlim", ibo_db=1.0, receiver="cnc")
    e3 = engine.sweep(snr[:3], None, 23, n_iter, seed=6, frame_id0=40)
    for p in range(3):
        ref = engine.simulate(23, n_iter, float(snr[p]), seed=6, frame_id0=40 + p * 23)["err"]
        assert np.array_equal(e3[p], ref), (p, e3[p], ref)


def test_radix64_and_radix16_frame_kernels_agree(monkeypatch):
    """The three forms of the frame kernel - the default scaled-tangent radix-64 kernel with its exchange token ring
    (frame_kernel_t64.cuh), the round-1 radix-64 kernel (MMIMO_FRAME_KERNEL=r64) and the three-pass radix-16 kernel (=r16) -
    are decompositions of the same frame: equalised samples agree to fp32 rounding, detections differ only on decision
    boundaries, and each kernel is bit-reproducible run to run (the four antenna groups are summed in a fixed order).
    The kernel form is read once per context (mmimo_create), hence one engine per form."""
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    engines = {}
    try:
        for kern in ("t64", "r64", "r16"):
            if kern == "t64":
                monkeypatch.delenv("MMIMO_FRAME_KERNEL", raising=False)
            else:
                monkeypatch.setenv("MMIMO_FRAME_KERNEL", kern)
            engines[kern] = LinkEngine(0)
        monkeypatch.delenv("MMIMO_FRAME_KERNEL", raising=False)
        # antenna counts that leave groups of the radix-64 kernels idle or ragged (1, 3, 5, 6) included: the token ring must
        # still be passed on by groups that have no antenna in a round
        for pa, n_ant, rx in (("softlim", 16, "mcnc"), ("rapp", 6, "mcnc"), ("softlim", 8, "cnc"), ("toi", 4, "mcnc"),
                              ("softlim", 1, "mcnc"), ("softlim", 5, "cnc"), ("none", 3, "mcnc"), ("softlim", 7, "mcnc")):
            outs = {}
            for kern, eng in engines.items():
                eng.configure(n_ant=n_ant, channel="los", pa=pa, ibo_db=1.0, receiver=rx)
                outs[kern] = eng.simulate(12, 3, 17.0, seed=31, dumps=("rx_eq", "det"))
                again = eng.simulate(12, 3, 17.0, seed=31, dumps=("rx_eq", "det"))
                assert np.array_equal(outs[kern]["err"], again["err"]) and np.array_equal(outs[kern]["rx_eq"], again["rx_eq"])
            b = outs["r16"]
            for kern in ("t64", "r64"):
                a = outs[kern]
                rel = np.linalg.norm(a["rx_eq"] - b["rx_eq"]) / np.linalg.norm(b["rx_eq"])
                assert rel < 2e-6, (kern, pa, rel)
                assert (a["det"] != b["det"]).mean() < 1e-3, (kern, pa)
                assert np.abs(a["err"].astype(np.int64) - b["err"].astype(np.int64)).max() <= 6, (kern, a["err"], b["err"])
            # the three forms are really three kernels: their samples differ in the last bits
            assert not np.array_equal(outs["t64"]["rx_eq"], outs["r16"]["rx_eq"]) or pa == "none"
    finally:
        for eng in engines.values():
            eng.close()


def test_cnc_loop_kernel_matches_single_kernel_form(monkeypatch):
    """CncReceiver frames: pass 0 in the frame kernel + the receiver loop in k_cnc_loop_t64 (one frame per 64-thread group,
    the default) against the whole frame in k_link_frame_t64 (MMIMO_CNC_SPLIT=0).  The loop kernel repeats the frame kernel's
    arithmetic operation by operation, so decisions, corrected symbols, per-frame errors and counters are IDENTICAL; frame
    counts that leave groups of the last CTA round without a frame (1, 2, 3 mod 4) and injected symbols are covered."""
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    engines = {}
    try:
        for split in ("1", "0"):
            monkeypatch.setenv("MMIMO_CNC_SPLIT", split)
            engines[split] = LinkEngine(0)
        monkeypatch.delenv("MMIMO_CNC_SPLIT", raising=False)
        rng = np.random.default_rng(5)
        for pa, n_ant, n_frames, n_iter, constel in (("softlim", 8, 13, 3, 64), ("rapp", 4, 602, 2, 16), ("toi", 2, 7, 5, 64),
                                                     ("softlim", 1, 4, 1, 4), ("none", 3, 9, 2, 256), ("softlim", 64, 150, 5, 64)):
            inject = None
            if n_frames == 13:
                inject = {"tx_idx": rng.integers(0, constel, (n_frames, 2048), dtype=np.uint8)}
            outs = {}
            for split, eng in engines.items():
                eng.configure(n_ant=n_ant, channel="los", pa=pa, ibo_db=0.5, receiver="cnc", constel_size=constel)
                kw = dict(seed=41, dumps=("rx_eq", "det", "corr", "err_per_frame"))
                if inject:
                    kw["inject"] = inject
                outs[split] = eng.simulate(n_frames, n_iter, 16.0, **kw)
            a, b = outs["1"], outs["0"]
            for key in ("err", "det", "err_per_frame", "rx_eq", "corr"):
                assert np.array_equal(a[key], b[key]), (pa, n_ant, n_frames, key)
            assert (a["err"][0] > 0 or constel == 4) and int(a["err_per_frame"].sum()) == int(a["err"].sum())
    finally:
        for eng in engines.values():
            eng.close()


@pytest.mark.parametrize("channel,csi_eps", [("two_path", 0.0), ("rayleigh", 0.0), ("los", 0.2), ("two_path", 0.1)])
def test_fused_norm_pass_is_bit_identical(engine, channel, csi_eps):
    """From 2 realisations per SM on, the kernel that writes the rows the precoder is derived from (k_gen_channel,
    k_gen_rayleigh, or k_csi_error for a channel with CSI error) also writes the MRT symbol scale (one CTA per realisation)
    and k_prep_realisation skips its norm pass; below that the realisation kernel computes it.  Same operation order, so a
    large batch and the same frames run as small batches give exactly the same counters and samples.  (A line-of-sight
    channel without CSI error goes through the closed-form k_prep_los at every batch size.)"""
    engine.configure(n_ant=16, channel=channel, pa="softlim", ibo_db=0.0, receiver="mcnc")
    n, n_iter = 320, 2                      # 320 >= 2 * 148: fused path
    big = engine.simulate(n, n_iter, 16.0, seed=9, frame_id0=500, csi_eps=csi_eps, dumps=("rx_eq",))
    err = np.zeros_like(big["err"])
    for c0 in range(0, n, 80):              # 80 < 2 * 148: norm pass in the realisation kernel
        part = engine.simulate(80, n_iter, 16.0, seed=9, frame_id0=500 + c0, csi_eps=csi_eps, dumps=("rx_eq",))
        err += part["err"]
        assert np.array_equal(part["rx_eq"], big["rx_eq"][c0:c0 + 80])
    assert np.array_equal(err, big["err"])
    assert err[0] > 0


def test_closed_form_los_realisation_matches_oracle(monkeypatch):
    """k_prep_los (closed form in the K path amplitudes, no pass over the channel rows) and k_prep_realisation
    (MMIMO_PREP_ANALYTIC=0) against the oracle's realisation bookkeeping (antenna_array.py:163-171,313-360, mp_model.py:290-329)
    for device-generated LOS channels at injected receiver positions; and both kernels give the same frames."""
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    cfg = lo.LinkConfig(n_ant=64, channel="los", ibo_db=1.5)
    rng = np.random.default_rng(12)
    pos = np.stack([lo.draw_rx_position(cfg, rng) for _ in range(4)])
    snr = float(lo.ebn0_to_snr(15.0, cfg.n_sub_carr, cfg.n_sub_carr, 64))
    outs = {}
    for knob in ("1", "0"):
        monkeypatch.setenv("MMIMO_PREP_ANALYTIC", knob)
        eng = LinkEngine(0)
        try:
            eng.configure(n_ant=64, channel="los", ibo_db=1.5)
            out = eng.simulate(4, 2, snr, seed=5, inject=dict(rx_pos=pos), dumps=("agc", "pa_p0", "alpha", "rx_eq", "det"))
        finally:
            eng.close()
        outs[knob] = out
        for i in range(4):
            real = lo.make_realisation(cfg, lo.draw_channel(cfg, pos[i]))
            assert rel_l2(out["agc"][i], real.ak_hk_vk_agc) < TOL, knob
            assert abs(out["pa_p0"][i] / real.pa[1] - 1) < TOL, knob
            assert np.allclose(out["alpha"][i], real.alpha_k, rtol=1e-5), knob
    a, b = outs["1"], outs["0"]
    assert rel_l2(a["rx_eq"], b["rx_eq"]) < 2e-6
    assert (a["det"] != b["det"]).mean() < 1e-3 and np.abs(a["err"].astype(np.int64) - b["err"].astype(np.int64)).max() <= 6


def test_pipelined_realisations_give_the_same_counters(monkeypatch):
    """mmimo_link_enqueue called back to back generates the channels of batch n + 1 on a second stream (and into a second
    buffer set) while the frame kernel of batch n runs (mmimo_ctx::PrepSet); the counters must equal those of the serial
    order (MMIMO_PIPELINE_PREP=0) exactly, with device-generated and with injected per-frame inputs."""
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    rng = np.random.default_rng(4)
    n, n_iter, snr = 300, 2, 21.0
    inj = []
    for b in range(4):
        pos = np.stack([(212.0 + rng.uniform(-5, 5), 212.0 + rng.uniform(-5, 5), 1.5) for _ in range(n)])
        inj.append(dict(rx_pos=pos, tx_idx=rng.integers(0, 64, (n, 2048), dtype=np.uint8),
                        noise=(rng.standard_normal((n, 2048)) + 1j * rng.standard_normal((n, 2048))).astype(np.complex64) * 1e-6))
    res = {}
    for knob in ("1", "0"):
        monkeypatch.setenv("MMIMO_PIPELINE_PREP", knob)
        eng = LinkEngine(0)
        try:
            out = []
            for chan in ("los", "two_path", "rayleigh"):
                eng.configure(n_ant=16, channel=chan, ibo_db=0.0)
                for b in range(4):
                    eng.enqueue(n, n_iter, snr, seed=3, frame_id0=b * n)
                out.append(eng.fetch(n_iter)[0].copy())
            eng.configure(n_ant=16, channel="los", ibo_db=0.0)
            for b in range(4):
                eng.enqueue(n, n_iter, snr, seed=3, frame_id0=b * n, inject=inj[b])
            out.append(eng.fetch(n_iter)[0].copy())
        finally:
            eng.close()
        res[knob] = out
    for a, b in zip(res["1"], res["0"]):
        assert np.array_equal(a, b), (a, b)
        assert a[0] > 0
