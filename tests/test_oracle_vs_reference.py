"""Live pin of oracle/link_oracle.py against the UNMODIFIED reference (skipped where the checkout is absent, e.g. on
the GPU box - the committed tests/golden/*.npz carry the same check there).  Small case: K = 2, one frame."""
import ctypes
import multiprocessing as mp

import numpy as np
import pytest

from oracle import link_oracle as lo
from oracle import refharness

pytestmark = pytest.mark.skipif(not refharness.reference_available(), reason="reference checkout not present")


@pytest.mark.parametrize("is_mcnc", [True, False])
def test_reference_link_simulate_equals_oracle(is_mcnc):
    import copy
    R = refharness.import_reference()
    modulation, distortion, transceiver, antenna_array, channel, noise, mp_model, utilities = (
        R[k] for k in ("modulation", "distortion", "transceiver", "antenna_array", "channel", "noise", "mp_model", "utilities"))
    n_ant, ibo, ebn0, iters, seeds = 2, 1.0, 14.0, np.array([0, 1, 2]), [5, 6, 7]
    mod = modulation.OfdmQamModem(constel_size=64, n_fft=4096, n_sub_carr=2048, cp_len=128)
    dist = distortion.SoftLimiter(0.0, mod.avg_sample_power)
    tx = transceiver.Transceiver(modem=copy.deepcopy(mod), impairment=copy.deepcopy(dist), center_freq=int(3.5e9), carrier_spacing=int(15e3))
    rx = transceiver.Transceiver(modem=copy.deepcopy(mod), impairment=copy.deepcopy(dist), cord_x=212, cord_y=212, cord_z=1.5,
                                 center_freq=int(3.5e9), carrier_spacing=int(15e3))
    arr = antenna_array.LinearArray(n_elements=n_ant, base_transceiver=tx, center_freq=int(3.5e9), wav_len_spacing=0.5, cord_z=15)
    ch = channel.MisoLosFd()
    ch.calc_channel_mat(tx_transceivers=arr.array_elements, rx_transceiver=rx, skip_attenuation=False)
    link = mp_model.Link(mod_obj=mod, array_obj=arr, std_rx_obj=rx, chan_obj=ch, noise_obj=noise.Awgn(snr_db=10, seed=1234),
                         rx_loc_var=10.0, n_err_min=int(1e12), bits_sent_max=mod.n_bits_per_ofdm_sym * 2, is_mcnc=is_mcnc)
    link.update_distortion(ibo_val_db=ibo)
    snr = float(utilities.ebn0_to_snr(ebn0, 2048, 2048, 64))
    link.set_snr(snr_db_val=snr)
    e = mp.Array(ctypes.c_double, 3, lock=True)
    b = mp.Array(ctypes.c_double, 3, lock=True)
    link.simulate(False, True, iters, seeds, e, b)
    cfg = lo.LinkConfig(n_ant=n_ant, ibo_db=ibo, channel="los")
    errs, bits = lo.link_simulate(cfg, snr, 2, seeds, 2, is_mcnc=is_mcnc)
    assert np.array_equal(errs, np.array(list(e)).astype(np.int64))
    assert np.array_equal(bits, np.array(list(b)).astype(np.int64))
