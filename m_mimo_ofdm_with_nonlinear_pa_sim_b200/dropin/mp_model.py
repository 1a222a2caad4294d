"""Drop-in for the reference's mp_model.py: `Link`, the Monte-Carlo frame loop.

`Link(...)`, `.simulate(...)`, `.update_distortion`, `.set_snr`, `.set_precoding_and_recalculate_agc` and
`.recalculate_agc` keep the reference's signatures and attributes (mp_model.py:32-329).  `simulate` runs whole
batches of frames in the FUSED CUDA frame kernel (mmimo_link_simulate) whenever the link is one the kernel
covers - n_fft 4096 / 2048 sub-carriers, one user, LOS / two-path / Rayleigh channel, soft-limiter or Rapp PA,
CNC or MCNC receiver, optional CSI error - and otherwise composes the same loop from the drop-in building blocks
(stage kernels).  Two random-number modes (attribute `rng_mode`, default from env MMIMO_RNG):

  "numpy"  (default) bits, receiver positions, Rayleigh coefficients, CSI-error and AWGN draws come from the
           reference's numpy PCG64 streams in the reference's order and are injected into the kernel: error
           counts are those the reference produces for the same seeds (up to fp32-vs-fp64 slicer ties).
  "philox" everything is drawn on the device (Philox4x32-10 keyed by seed and global frame index): full speed,
           statistically equivalent results.

The stopping rule (n_err_min / bits_sent_max per iteration slot, mp_model.py:181-187) is applied frame by
frame to the per-frame error counts the kernel returns, so single-worker results do not depend on the batch
size.  The caller-owned shared `multiprocessing.Array` counters are updated like the reference does."""
import copy
import os

import numpy as np
from numpy import ndarray

import channel
import corrector
import distortion
import modulation
import utilities
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi

_BATCH_MAX = 256


class Link():
    def __init__(self, mod_obj, array_obj, std_rx_obj, chan_obj, noise_obj, rx_loc_var: float, n_err_min: int,
                 bits_sent_max: int, is_mcnc: bool = False, csi_epsylon: float = None):
        self.my_mod = copy.deepcopy(mod_obj)
        self.my_array = copy.deepcopy(array_obj)
        self.my_standard_rx = copy.deepcopy(std_rx_obj)
        self.rx_loc_x = self.my_standard_rx.cord_x
        self.rx_loc_y = self.my_standard_rx.cord_y
        self.my_noise = copy.deepcopy(noise_obj)
        self.my_csi_noise = copy.deepcopy(noise_obj)
        self.csi_epsylon = csi_epsylon
        if isinstance(chan_obj, channel.MisoQuadrigaFd):
            self.is_quadriga = True
            self.channel_model_str = chan_obj.channel_model_str
            los = channel.MisoLosFd()
            los.calc_channel_mat(tx_transceivers=array_obj.array_elements, rx_transceiver=std_rx_obj, skip_attenuation=False)
            self.my_miso_chan = los
            if self.csi_epsylon is not None:
                self.my_miso_chan_csi_err = los
        else:
            self.is_quadriga = False
            self.my_miso_chan = copy.deepcopy(chan_obj)
            if self.csi_epsylon is not None:
                self.my_miso_chan_csi_err = copy.deepcopy(self.my_miso_chan)
        if is_mcnc:
            rx_chan = self.my_miso_chan if self.csi_epsylon is None else self.my_miso_chan_csi_err
            self.my_cnc_rx = corrector.McncReceiver(self.my_array, rx_chan)
        else:
            self.my_cnc_rx = corrector.CncReceiver(copy.deepcopy(array_obj.base_transceiver.modem),
                                                   copy.deepcopy(array_obj.base_transceiver.impairment))
        self.my_noise.rng_gen = np.random.default_rng(0)
        self.loc_rng = np.random.default_rng(1)
        self.bit_rng = np.random.default_rng(2)
        self.my_csi_noise.rng_gen = np.random.default_rng(3)
        self.rx_loc_var = rx_loc_var
        self.n_ant_val = len(self.my_array.array_elements)
        self.n_bits_per_ofdm_sym = self.my_mod.n_bits_per_ofdm_sym
        self.n_sub_carr = self.my_mod.n_sub_carr
        self.ibo_val_db = self.my_array.array_elements[0].impairment.ibo_db
        self.n_err_min = n_err_min
        self.bits_sent_max = bits_sent_max
        self.rng_mode = os.environ.get("MMIMO_RNG", "numpy")
        self.philox_seed = None  # "philox" mode: defaults to seed_arr[0] of simulate()
        self._engine = None
        self._engine_pid = None
        self.set_precoding_and_recalculate_agc()

    # the CUDA context must not travel through fork / pickle / deepcopy
    def __getstate__(self):
        st = dict(self.__dict__)
        st["_engine"] = None
        st["_engine_pid"] = None
        return st

    # ---- bookkeeping with the reference's attribute names -------------------------------------------
    def update_distortion(self, ibo_val_db: float) -> None:
        self.my_array.update_distortion(ibo_db=ibo_val_db, avg_sample_pow=self.my_mod.avg_sample_power)
        if isinstance(self.my_cnc_rx, corrector.CncReceiver):
            self.my_cnc_rx.update_distortion(ibo_db=ibo_val_db)
        self.ibo_val_db = self.my_array.array_elements[0].impairment.ibo_db
        self.recalculate_agc(ak_part_only=True)

    def set_snr(self, snr_db_val: float) -> None:
        self.my_noise.snr_db = snr_db_val

    def set_precoding_and_recalculate_agc(self) -> None:
        """MRT precoding from the (possibly CSI-error-perturbed) channel + AGC vectors (mp_model.py:253-288)."""
        if self.csi_epsylon is None:
            self.my_array.set_precoding_matrix(channel_mat_fd=self.my_miso_chan.channel_mat_fd, mr_precoding=True)
            self.recalculate_agc(channel_mat_fd=self.my_miso_chan.channel_mat_fd)
            return
        n_sc, n_fft = self.my_mod.n_sub_carr, self.my_mod.n_fft
        h = np.copy(self.my_miso_chan.channel_mat_fd)
        bins = np.concatenate((np.arange(n_fft - n_sc // 2, n_fft), np.arange(1, n_sc // 2 + 1)))
        for k in range(h.shape[0]):
            sc = h[k, bins]
            p = np.sum(np.abs(sc) ** 2) / len(sc)
            nz = self.my_noise.rng_gen.standard_normal((len(sc), 2)).view(np.complex128)[:, 0]
            h[k, bins] = _abi.get_ops().awgn_add(np.sqrt(1 - np.power(self.csi_epsylon, 2)) * sc, nz,
                                                 0.5 * np.sqrt(2 * p) * self.csi_epsylon)
        self.my_miso_chan_csi_err.channel_mat_fd = h
        self.my_array.set_precoding_matrix(channel_mat_fd=h, mr_precoding=True)
        self.recalculate_agc(channel_mat_fd=h)

    def recalculate_agc(self, channel_mat_fd: ndarray = None, ak_part_only: bool = False) -> None:
        """hk_vk / ak_hk_vk AGC vectors and noise scalers (mp_model.py:290-329)."""
        if not ak_part_only:
            self._agc_chan = channel_mat_fd
            self.my_array.update_distortion(ibo_db=self.ibo_val_db, avg_sample_pow=self.my_mod.avg_sample_power)
            if isinstance(self.my_cnc_rx, corrector.CncReceiver):
                self.my_cnc_rx.update_distortion(ibo_db=self.ibo_val_db)
        vk_mat = self.my_array.get_precoding_mat()
        res = _abi.get_ops().agc(self._agc_chan, vk_mat, 0, self.ibo_val_db, self.my_mod.n_sub_carr)
        self.vk_pow_vec = res["vk_pow"]
        if not ak_part_only:
            self.hk_vk_noise_scaler = float(res["scalers"][1])
            self.hk_vk_agc_nfft = res["hkvk_nfft"]
        self.ak_hk_vk_noise_scaler = float(res["scalers"][0])
        self.ak_hk_vk_agc_nfft = res["agc_nfft"]
        if isinstance(self.my_cnc_rx, corrector.McncReceiver):
            self.my_cnc_rx.agc_corr_vec = self.ak_hk_vk_agc_nfft

    # ---- simulate ------------------------------------------------------------------------------------
    def simulate(self, incl_clean_run: bool, reroll_chan: bool, cnc_n_iter_lst: list, seed_arr: list,
                 n_err_shared_arr, n_bits_sent_shared_arr) -> None:
        """Monte-Carlo frame loop of one worker (mp_model.py:89-222): frames are simulated until every result slot
        (optional clean run + one slot per entry of cnc_n_iter_lst) has n_err_min errors or bits_sent_max bits in
        the shared counters."""
        if self.is_quadriga:
            self.my_miso_chan = channel.MisoQuadrigaFd(tx_transceivers=self.my_array.array_elements,
                                                       rx_transceiver=self.my_standard_rx,
                                                       channel_model_str=self.channel_model_str)
            if self.csi_epsylon is not None:
                self.my_miso_chan_csi_err = channel.MisoQuadrigaFd(tx_transceivers=self.my_array.array_elements,
                                                                   rx_transceiver=self.my_standard_rx,
                                                                   channel_model_str=self.channel_model_str,
                                                                   start_matlab_eng=False)
        if isinstance(self.my_cnc_rx, corrector.McncReceiver):
            self.my_cnc_rx.channel = self.my_miso_chan if self.csi_epsylon is None else self.my_miso_chan_csi_err
        self.bit_rng = np.random.default_rng(seed_arr[0])
        self.my_noise.rng_gen = np.random.default_rng(seed_arr[1])
        self.loc_rng = np.random.default_rng(seed_arr[2])
        if self.csi_epsylon is not None:
            self.my_csi_noise.rng_gen = np.random.default_rng(seed_arr[3])
        iters = np.asarray(cnc_n_iter_lst).astype(np.int64).reshape(-1)
        if self._fused_ok(bool(reroll_chan)) and (iters.size == 0 or int(iters.max()) <= 62):  # the fused kernel runs up to 62 iterations
            self._simulate_fused(bool(incl_clean_run), bool(reroll_chan), iters, seed_arr, n_err_shared_arr, n_bits_sent_shared_arr)
        else:
            self._simulate_composed(bool(incl_clean_run), bool(reroll_chan), iters, n_err_shared_arr, n_bits_sent_shared_arr)

    # -- building-block path: any shape / channel the stage kernels cover ---------------------------------
    def _reroll(self):
        if not isinstance(self.my_miso_chan, channel.MisoRayleighFd):
            half = self.rx_loc_var / 2.0
            # like the reference, BOTH offsets are added to rx_loc_x (mp_model.py:144-147, :194-197)
            self.my_standard_rx.set_position(cord_x=self.rx_loc_x + self.loc_rng.uniform(low=-half, high=half),
                                             cord_y=self.rx_loc_x + self.loc_rng.uniform(low=-half, high=half),
                                             cord_z=self.my_standard_rx.cord_z)
            self.my_miso_chan.calc_channel_mat(tx_transceivers=self.my_array.array_elements,
                                               rx_transceiver=self.my_standard_rx)
        else:
            self.my_miso_chan.reroll_channel_coeffs()
        self.set_precoding_and_recalculate_agc()

    def _simulate_composed(self, incl_clean_run, reroll_chan, iters, err_arr, bits_arr):
        res_idx = 0
        if incl_clean_run:
            res_idx = 1
            while err_arr[0] < self.n_err_min and bits_arr[0] < self.bits_sent_max:
                if reroll_chan:
                    self._reroll()
                tx_bits = self.bit_rng.choice((0, 1), self.n_bits_per_ofdm_sym)
                sig = self.my_array.transmit(tx_bits, out_domain_fd=True, return_both=False, skip_dist=True)
                sig = self.my_miso_chan.propagate(in_sig_mat=sig)
                sig = self.my_noise.process(sig, avg_sample_pow=self.my_mod.avg_symbol_power * self.hk_vk_noise_scaler)
                sig = np.divide(sig, self.hk_vk_agc_nfft)
                td = utilities.to_time_domain(sig)
                rx_bits = self.my_standard_rx.receive(np.concatenate((td[-self.my_mod.cp_len:], td)))
                err_arr[0] += utilities.count_mismatched_bits(tx_bits, rx_bits)
                bits_arr[0] += self.my_mod.n_bits_per_ofdm_sym
        err_np = np.frombuffer(err_arr.get_obj())
        bits_np = np.frombuffer(bits_arr.get_obj())
        while True:
            use = np.logical_and(err_np[res_idx:] < self.n_err_min, bits_np[res_idx:] < self.bits_sent_max)
            if not use.any():
                break
            if reroll_chan:
                self._reroll()
            tx_bits = self.bit_rng.choice((0, 1), self.n_bits_per_ofdm_sym)
            sig = self.my_array.transmit(tx_bits, out_domain_fd=True, skip_dist=False)
            sig = self.my_miso_chan.propagate(in_sig_mat=sig)
            sig = self.my_noise.process(sig, avg_sample_pow=self.my_mod.avg_symbol_power * self.ak_hk_vk_noise_scaler)
            sig = np.divide(sig, self.ak_hk_vk_agc_nfft)
            rx_bits_lst = self.my_cnc_rx.receive(n_iters_lst=iters[use], in_sig_fd=sig)
            order = np.argsort(iters[use], kind="stable")  # receive() returns ascending iteration order
            slots = (np.nonzero(use)[0] + res_idx)[order]
            log = getattr(self, "frame_log", None)  # a list: per-frame decisions are recorded there (parity tests)
            if log is not None:
                log.append(dict(iters=np.sort(iters[use]), bits=[np.asarray(b) for b in rx_bits_lst], tx_bits=np.asarray(tx_bits)))
            for slot, rx_bits in zip(slots, rx_bits_lst):
                err_arr[slot] += utilities.count_mismatched_bits(tx_bits, rx_bits)
                bits_arr[slot] += self.n_bits_per_ofdm_sym

    # -- fused path: batches of frames through mmimo_link_simulate ------------------------------------------
    def _fused_ok(self, reroll_chan: bool = True) -> bool:
        if os.environ.get("MMIMO_LINK_PATH", "fused") != "fused" or self.is_quadriga:
            return False
        m = self.my_mod
        if (m.n_fft, m.n_sub_carr, m.n_users) != (4096, 2048, 1) or m.constel_size not in (4, 16, 64, 256):
            return False
        if type(self.my_miso_chan) not in (channel.MisoLosFd, channel.MisoTwoPathFd, channel.MisoRayleighFd):
            # any other channel object (MisoRandomPathsFd, a user's own class) is carried by its channel_mat_fd: possible
            # when the channel stays fixed; re-rolling needs calc_channel_mat / reroll_channel_coeffs of a known type
            if reroll_chan or not hasattr(self.my_miso_chan, "channel_mat_fd"):
                return False
        arr = self.my_array
        if arr.n_elements > 256 or arr._uniform() is None:
            return False
        e0 = arr.array_elements[0]
        if type(e0.impairment) not in (distortion.SoftLimiter, distortion.Rapp):
            return False
        if modulation._table_scale(m._constellation) != 1.0 or modulation._table_scale(e0.modem._constellation) != 1.0:
            return False  # a corrected (alpha-scaled) constellation goes through the building-block path
        if len({t.tx_ant_gain_db for t in arr.array_elements}) != 1:
            return False
        if type(self.my_cnc_rx) not in (corrector.McncReceiver, corrector.CncReceiver):
            return False
        return True

    def _get_engine(self, injected_channel: bool):
        from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
        if self._engine is None or self._engine_pid != os.getpid():
            dev = os.environ.get("MMIMO_DEVICE", os.environ.get("LOCAL_RANK"))
            self._engine = LinkEngine(int(dev) if dev is not None else -1)
            self._engine_pid = os.getpid()
        arr, rx, m = self.my_array, self.my_standard_rx, self.my_mod
        e0 = arr.array_elements[0]
        kind = {channel.MisoLosFd: "los", channel.MisoTwoPathFd: "two_path",
                channel.MisoRayleighFd: "rayleigh"}.get(type(self.my_miso_chan), "injected")
        pa = "rapp" if isinstance(e0.impairment, distortion.Rapp) else "softlim"
        tx_pos = np.array([[t.cord_x, t.cord_y, t.cord_z] for t in arr.array_elements], dtype=np.float64)
        self._engine.configure(constel_size=m.constel_size, n_ant=arr.n_elements, pa=pa, ibo_db=float(self.ibo_val_db),
                               p_hardness=float(getattr(e0.impairment, "p_hardness", 0.0)),
                               channel="injected" if injected_channel else kind,
                               receiver="mcnc" if isinstance(self.my_cnc_rx, corrector.McncReceiver) else "cnc",
                               center_freq=float(rx.center_freq), carrier_spacing=float(rx.carrier_spacing),
                               rx_pos=(float(self.rx_loc_x), float(self.rx_loc_y), float(rx.cord_z)),
                               rx_loc_var=float(self.rx_loc_var), ant_gain_db=float(e0.tx_ant_gain_db + rx.rx_ant_gain_db),
                               tx_pos=tx_pos, freqs=channel._freqs(rx))
        return self._engine

    def _rngs(self):
        gens = [self.bit_rng, self.my_noise.rng_gen, self.loc_rng]
        if hasattr(self.my_miso_chan, "rng_gen"):
            gens.append(self.my_miso_chan.rng_gen)
        return gens

    def _draw_batch(self, n, reroll_chan, prefix):
        """Host inputs of `n` frames drawn from the reference's numpy streams in the reference's per-frame order
        (mp_model.py:190-212): position or Rayleigh coefficients, CSI-error noise, bits, AWGN.  Also returns the
        generator states after every frame so that an early stop can rewind to exactly the frames it used."""
        m, K = self.my_mod, self.n_ant_val
        n_fft, n_sc, bps = m.n_fft, m.n_sub_carr, m.n_bits_per_symbol
        bins = np.concatenate((np.arange(n_fft - n_sc // 2, n_fft), np.arange(1, n_sc // 2 + 1)))
        rayleigh = isinstance(self.my_miso_chan, channel.MisoRayleighFd)
        half = self.rx_loc_var / 2.0
        w = (1 << np.arange(bps - 1, -1, -1)).astype(np.int64)
        inj = {prefix + "tx_idx": np.empty((n, n_sc), dtype=np.uint8), prefix + "noise": np.empty((n, n_sc), dtype=np.complex64)}
        if reroll_chan and not rayleigh:
            inj[prefix + "rx_pos"] = np.empty((n, 3), dtype=np.float64)
        if reroll_chan and rayleigh:
            inj[prefix + "h_true"] = np.empty((n, K, n_sc), dtype=np.complex64)
            att = self.my_miso_chan.los_fd_att_mat[:, bins]
        if self.csi_epsylon is not None and reroll_chan:
            inj[prefix + "csi_noise"] = np.empty((n, K, n_sc), dtype=np.complex64)
        gens = self._rngs()
        states = [[g.bit_generator.state for g in gens]]  # [0] = before the batch, [f + 1] = after frame f
        for f in range(n):
            if reroll_chan:
                if not rayleigh:
                    inj[prefix + "rx_pos"][f] = (self.rx_loc_x + self.loc_rng.uniform(low=-half, high=half),
                                                 self.rx_loc_x + self.loc_rng.uniform(low=-half, high=half),
                                                 self.my_standard_rx.cord_z)
                else:
                    co = self.my_miso_chan.rng_gen.standard_normal(size=(K, n_fft * 2)).view(dtype=np.complex128) / np.sqrt(2.0)
                    inj[prefix + "h_true"][f] = co[:, bins] * att
                if self.csi_epsylon is not None:
                    for k in range(K):
                        inj[prefix + "csi_noise"][f, k] = self.my_noise.rng_gen.standard_normal((n_sc, 2)).view(np.complex128)[:, 0]
            bits = self.bit_rng.choice((0, 1), self.n_bits_per_ofdm_sym)
            inj[prefix + "tx_idx"][f] = bits.reshape(n_sc, bps) @ w
            inj[prefix + "noise"][f] = self.my_noise.rng_gen.standard_normal((n_fft, 2)).view(np.complex128)[bins, 0]
            states.append([g.bit_generator.state for g in gens])
        if not reroll_chan:  # the channel and the channel knowledge stay what the object holds now
            h = np.asarray(self.my_miso_chan.channel_mat_fd)[:, bins].astype(np.complex64)
            inj[prefix + "h_true"] = np.broadcast_to(h, (n, K, n_sc))
            if self.csi_epsylon is not None:
                hl = np.asarray(self.my_miso_chan_csi_err.channel_mat_fd)[:, bins].astype(np.complex64)
                inj[prefix + "h_loop"] = np.broadcast_to(hl, (n, K, n_sc))
        return inj, gens, states

    def _batch_size(self, bits_now):
        left = int(np.ceil((self.bits_sent_max - float(np.min(bits_now))) / self.n_bits_per_ofdm_sym))
        cap = _BATCH_MAX if self.rng_mode != "philox" else 4096
        if self.rng_mode != "philox" and (isinstance(self.my_miso_chan, channel.MisoRayleighFd) or self.csi_epsylon is not None):
            cap = max(1, min(cap, (1 << 27) // (self.n_ant_val * self.n_sub_carr * 8)))  # <= 128 MiB of injected matrices
        return max(1, min(cap, left))

    def _simulate_fused(self, incl_clean_run, reroll_chan, iters, seed_arr, err_arr, bits_arr):
        philox = self.rng_mode == "philox"
        eng = self._get_engine(injected_channel=not reroll_chan)
        snr = float(self.my_noise.snr_db)
        eps = float(self.csi_epsylon) if (self.csi_epsylon is not None and reroll_chan) else 0.0
        B = self.n_bits_per_ofdm_sym
        seed = int(self.philox_seed if self.philox_seed is not None else seed_arr[0])
        frame_id = 0
        res_idx = 0
        if incl_clean_run:
            res_idx = 1
            while err_arr[0] < self.n_err_min and bits_arr[0] < self.bits_sent_max:
                n = self._batch_size([bits_arr[0]])
                if philox:
                    inj, gens, states = ({} if reroll_chan else self._fixed_channel(n, "clean_")), None, None
                else:
                    inj, gens, states = self._draw_batch(n, reroll_chan, "clean_")
                out = eng.simulate(n, 0, snr, seed=seed, frame_id0=frame_id, incl_clean_run="only", csi_eps=eps, inject=inj,
                                   dumps=("clean_err_per_frame",), noise_unit=not philox)
                frame_id += n
                used = 0
                for f in range(n):
                    if not (err_arr[0] < self.n_err_min and bits_arr[0] < self.bits_sent_max):
                        break
                    err_arr[0] += int(out["clean_err_per_frame"][f])
                    bits_arr[0] += B
                    used += 1
                self._rewind(gens, states, used, n)
        err_np = np.frombuffer(err_arr.get_obj())
        bits_np = np.frombuffer(bits_arr.get_obj())
        while True:
            use = np.logical_and(err_np[res_idx:] < self.n_err_min, bits_np[res_idx:] < self.bits_sent_max)
            if not use.any():
                break
            n_iter = int(np.max(iters[use]))
            n = self._batch_size(bits_np[res_idx:][use])
            if philox:
                inj, gens, states = ({} if reroll_chan else self._fixed_channel(n, "")), None, None
            else:
                inj, gens, states = self._draw_batch(n, reroll_chan, "")
            log = getattr(self, "frame_log", None)  # a list: per-frame decisions are recorded there (parity tests)
            out = eng.simulate(n, n_iter, snr, seed=seed, frame_id0=(1 << 40) + frame_id, csi_eps=eps, inject=inj,
                               dumps=("err_per_frame", "det") if log is not None else ("err_per_frame",), noise_unit=not philox)
            frame_id += n
            epf = out["err_per_frame"]
            used = 0
            for f in range(n):
                use = np.logical_and(err_np[res_idx:] < self.n_err_min, bits_np[res_idx:] < self.bits_sent_max)
                if not use.any():
                    break
                for j in np.nonzero(use)[0]:
                    if iters[j] <= n_iter:
                        err_arr[res_idx + j] += int(epf[f, iters[j]])
                        bits_arr[res_idx + j] += B
                if log is not None:
                    log.append(dict(det=np.array(out["det"][f]), err=np.array(epf[f]), n_iter=n_iter))
                used += 1
            self._rewind(gens, states, used, n)
            last = (inj, used - 1) if used > 0 else None
        if reroll_chan and not philox and 'last' in locals() and last is not None:
            self._leave_state_of(*last)

    def _leave_state_of(self, inj, f):
        """The reference leaves the objects in the state of the LAST simulated realisation (receiver position, channel
        matrix, precoder, AGC vectors, McncReceiver.agc_corr_vec); scripts that read them after simulate(), or call
        simulate(reroll_chan=False) next, see the same.  LOS / two-path channels in "numpy" RNG mode: re-create that
        realisation through the building blocks (one channel + precoding + AGC evaluation per simulate() call).  Rayleigh
        and "philox" mode keep the pre-call state."""
        if "rx_pos" not in inj or isinstance(self.my_miso_chan, channel.MisoRayleighFd) or self.csi_epsylon is not None:
            return  # (re-deriving a CSI-error channel would draw from my_noise.rng_gen and shift the noise stream)
        x, y, z = (float(v) for v in inj["rx_pos"][f])
        self.my_standard_rx.set_position(cord_x=x, cord_y=y, cord_z=z)
        self.my_miso_chan.calc_channel_mat(tx_transceivers=self.my_array.array_elements, rx_transceiver=self.my_standard_rx)
        self.set_precoding_and_recalculate_agc()

    def _fixed_channel(self, n, prefix):
        m = self.my_mod
        bins = np.concatenate((np.arange(m.n_fft - m.n_sub_carr // 2, m.n_fft), np.arange(1, m.n_sub_carr // 2 + 1)))
        h = np.asarray(self.my_miso_chan.channel_mat_fd)[:, bins].astype(np.complex64)
        inj = {prefix + "h_true": np.broadcast_to(h, (n,) + h.shape)}
        if self.csi_epsylon is not None:
            hl = np.asarray(self.my_miso_chan_csi_err.channel_mat_fd)[:, bins].astype(np.complex64)
            inj[prefix + "h_loop"] = np.broadcast_to(hl, (n,) + hl.shape)
        return inj

    @staticmethod
    def _rewind(gens, states, used, n):
        """Frames drawn beyond the stop point must not have consumed the streams (the reference draws per frame)."""
        if gens is None or used >= n:
            return
        for g, st in zip(gens, states[used]):
            g.bit_generator.state = st


# One worker process per GPU context: CUDA does not survive fork(), and the reference's runners start their
# workers with multiprocessing's platform default (fork on Linux).  Importing this module therefore switches the
# default start method to "spawn" (set MMIMO_MP_START=keep to leave it alone).
if os.environ.get("MMIMO_MP_START", "spawn") != "keep":
    import multiprocessing as _mp
    try:
        _mp.set_start_method(os.environ.get("MMIMO_MP_START", "spawn"), force=True)
    except (RuntimeError, ValueError):
        pass
