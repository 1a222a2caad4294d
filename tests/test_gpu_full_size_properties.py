"""Size-independent properties of the fused path at BASELINE.json's full sizes (K = 64 ... 256 antennas, n_fft 4096, thousands
of frames, device RNG) - where the CPU oracle would take hours: round trips, sharding invariance (a checksum of
checksums), determinism, monotonicity, and the extreme configurations of the C ABI."""
import numpy as np
import pytest

from oracle import link_oracle as lo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    eng = LinkEngine(0)
    yield eng
    eng.close()


def snr_of(ebn0):
    return float(lo.ebn0_to_snr(ebn0, 2048, 2048, 64))


@pytest.mark.parametrize("channel", ["los", "two_path", "rayleigh"])
def test_round_trip_without_pa_is_error_free(engine, channel):
    """soft limiter backed off by 40 dB (never clips, alpha = 1), 60 dB SNR: bits -> QAM -> precode -> IFFT -> FFT -> channel
    -> AGC -> slicer returns every bit, every iteration"""
    engine.configure(n_ant=64, channel=channel, pa="softlim", ibo_db=40.0, receiver="mcnc")
    out = engine.simulate(512, 3, 60.0, seed=17, incl_clean_run=True)
    assert out["err"].tolist() == [0, 0, 0, 0] and int(out["clean_err"][0]) == 0
    assert int(np.ravel(out["bits"])[0]) == 512 * 12288


def test_counters_do_not_depend_on_sharding_or_batching(engine):
    """4096 frames at once == 4 shards with their global frame ids (what --gpus N relies on), and twice the same call
    gives the same counters"""
    engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=0.0, receiver="mcnc")
    snr = snr_of(15.0)
    whole = engine.simulate(4096, 5, snr, seed=2137)["err"]
    again = engine.simulate(4096, 5, snr, seed=2137)["err"]
    assert np.array_equal(whole, again)
    parts = sum(engine.simulate(1024, 5, snr, seed=2137, frame_id0=1024 * i)["err"].astype(np.int64) for i in range(4))
    assert np.array_equal(whole.astype(np.int64), parts)
    ber = whole / (4096 * 12288)
    assert np.all(np.diff(ber) < 0), ber           # every MCNC iteration helps at 15 dB, IBO 0 dB
    assert 0.07 < ber[0] < 0.11 and ber[5] < 0.01


def test_extreme_configurations(engine):
    """largest array, most iterations, every constellation, single antenna: run and stay sane"""
    engine.configure(n_ant=256, channel="los", pa="softlim", ibo_db=0.0, receiver="mcnc")
    out = engine.simulate(8, 62, snr_of(15.0), seed=3)
    ber = out["err"] / out["bits"]
    assert ber.shape == (63,) and ber[62] <= ber[0] and np.all(ber < 0.5)
    for constel in (4, 16, 256):
        engine.configure(constel_size=constel, n_ant=64, channel="los", pa="rapp", ibo_db=3.0, receiver="cnc")
        out = engine.simulate(16, 2, 40.0, seed=4)
        assert np.all(out["err"] <= out["bits"]) and int(np.ravel(out["bits"])[0]) == 16 * 2048 * int(np.log2(constel))
    engine.configure(n_ant=1, channel="los", pa="softlim", ibo_db=0.0, receiver="cnc")
    out = engine.simulate(64, 5, snr_of(25.0), seed=5)
    assert out["err"][5] < out["err"][0]


def test_two_user_and_sweep_at_full_size(engine):
    engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=0.0, receiver="mcnc")
    snr = snr_of(15.0)
    a = engine.mu_simulate(512, 3, snr, precoding="zf", seed=8)
    b = sum(engine.mu_simulate(128, 3, snr, precoding="zf", seed=8, frame_id0=128 * i)["err"].astype(np.int64) for i in range(4))
    assert np.array_equal(a["err"].astype(np.int64), b)
    # PA backed off by 40 dB, no noise: zero forcing separates the users perfectly
    engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=40.0, receiver="mcnc")
    z = engine.mu_simulate(64, 1, 80.0, precoding="zf", seed=9)
    assert z["err"].tolist() == [[0, 0], [0, 0]]
    # a sweep over 27 Eb/N0 points x 64 frames: BER falls monotonically with Eb/N0 for the last iteration
    engine.configure(n_ant=64, channel="los", pa="softlim", ibo_db=3.0, receiver="mcnc")
    e = engine.sweep([snr_of(v) for v in np.arange(10.0, 23.1, 0.5)], None, 64, 4, seed=10)
    last = e[:, 4].astype(np.int64)
    assert last[0] > last[13] > last[26] and np.all(np.diff(last[::4]) <= 0)
