// Signal-to-distortion evaluator: the per-snapshot powers that
// main_beampatterns_plotting/sdr_mp.py:124-151 computes from AntennaArray.transmit(return_both=True) and
// Miso*Fd.propagate(sum_signals=False):
//     desired    = sum_s | sum_k alpha_k h_k[s] X_clean_k[s] |^2
//     distortion = sum_s | sum_k h_k[s] (X_k[s] - alpha_k X_clean_k[s]) |^2
// fused like the frame kernel: one CTA per snapshot, antennas looped, both antenna sums in registers, the [K, n_fft]
// signal matrices of the reference are never materialised.
#pragma once
#include "frame_kernel.cuh"

namespace mmimo {

// The evaluator itself is the SDR epilogue of k_link_frame_t64 (frame_kernel_t64.cuh, FrameArgs::sdr_desired): with MRT
// from the true channel the desired component of sub-carrier s is sym[s] * agc[s], agc = sum_k alpha_k h_k v_k being the
// AGC vector the realisation kernels already produce, so a snapshot is pass 0 of the frame kernel without noise followed
// by two reductions.  (Round 1 had a separate radix-16 kernel here that accumulated both antenna sums itself.)
}  // namespace mmimo
