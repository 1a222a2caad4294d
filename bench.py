#!/usr/bin/env python
"""bench.py - 64-antenna LOS MCNC (I=5) OFDM frames/s on B200 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W]           # this repo's CUDA path
  python bench.py --impl reference [--gpus N ...]           # the reference algorithm on the host CPUs

One "step" = one batch of --frames Monte-Carlo frames per GPU through the fused link kernels
(channel generation -> MRT precoder/AGC -> frame kernel), followed by the counter exchange.
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for how every field is obtained.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_FFT, N_SC, CONSTEL = 4096, 2048, 64
METRIC = "64-ant MCNC OFDM frames/s"


def flops_per_frame(n_ant, n_iter):
    """Algorithmic FFT work of one frame: 2*K*(I+1) transforms of 5*N*log2(N) flops."""
    return 2.0 * n_ant * (n_iter + 1) * 5.0 * N_FFT * math.log2(N_FFT)


def snr_db_from_ebn0(ebn0_db):
    # utilities.ebn0_to_snr with n_fft := n_sub_carr as the scripts call it (main_mp_miso_mcnc_ber_vs_ebn0.py:114)
    return 10 * math.log10(10 ** (ebn0_db / 10) * math.log2(CONSTEL))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
                except ValueError:
                    continue
                for n, v in zip(names, f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                       samples=len(sm), power_w_max=float(max(pw)))
        return out


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


# ------------------------------------------------------------------------------------------------
# CPU legs: the oracle port of the reference algorithm (numpy complex128), the only code that may
# execute oracle/ here.
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    n_ant, n_iter, ebn0, n_frames, seed = args
    for v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[v] = "1"
    from oracle import link_oracle as lo
    cfg = lo.LinkConfig(n_ant=n_ant, channel="los", ibo_db=0.0)
    snr = float(lo.ebn0_to_snr(ebn0, cfg.n_sub_carr, cfg.n_sub_carr, CONSTEL))
    t0 = time.perf_counter()
    errs, bits = lo.link_simulate(cfg, snr, n_iter, [seed, seed + 1, seed + 2], n_frames)
    return time.perf_counter() - t0, errs.tolist(), bits.tolist()


def cpu_baseline_single(n_ant, n_iter, ebn0, budget_s=12.0):
    """Oracle port, 1 core, bounded sample of the same workload."""
    dt, _, _ = _cpu_worker((n_ant, n_iter, ebn0, 2, 11))  # warm-up + rate estimate
    n = int(max(4, min(400, budget_s / max(dt / 2, 1e-4))))
    dt, errs, bits = _cpu_worker((n_ant, n_iter, ebn0, n, 21))
    return {"value": n / dt, "unit": "frames/s", "cores": 1, "kind": "port",
            "sample": "%d frames of the bench workload (K=%d LOS MRT, soft limiter IBO 0 dB, Eb/N0 %.0f dB, MCNC I=%d) "
                      "through oracle/link_oracle.py (numpy complex128, pocketfft), 1 process" % (n, n_ant, ebn0, n_iter),
            "ber": [e / b for e, b in zip(errs, bits)]}


def run_reference(a):
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        per = max(1, a.ref_frames_per_core)
        jobs = lambda s: [(a.n_ant, a.n_iter, a.ebn0, per, 1000 * s + 7 * i) for i in range(cores)]
        for w in range(a.warmup):
            pool.map(_cpu_worker, [(a.n_ant, a.n_iter, a.ebn0, 1, i) for i in range(cores)])
        t0 = time.perf_counter()
        for s in range(a.steps):
            pool.map(_cpu_worker, jobs(s))
        dt = time.perf_counter() - t0
    frames = a.steps * cores * per
    val = frames / dt
    sample = ("%d frames per step (%d per core x %d cores) of the bench workload through oracle/link_oracle.py "
              "(numpy port of the reference algorithm; the reference itself is Python and cannot travel to this box)"
              % (cores * per, per, cores))
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "c128", "data": "synthetic",
            "config": workload_config(a, cores * per),
            "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample,
                             "reference_measured": reference_measured()},
            "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(a, frames_per_step):
    return {"workload": "configs[1]: MISO LOS %d-antenna ULA, MRT precoding, soft limiter IBO 0 dB, 64-QAM, n_fft 4096 / "
                        "2048 sub-carriers, Eb/N0 %.0f dB, MCNC %d iterations, receiver position re-drawn per frame"
                        % (a.n_ant, a.ebn0, a.n_iter),
            "n_ant": a.n_ant, "n_iter": a.n_iter, "ebn0_db": a.ebn0, "frames_per_step_per_gpu": frames_per_step,
            "sharding": "frames by global frame id, counters all-reduced",
            "cache": "inputs larger than L2: every step streams %.1f GiB of per-frame channel matrices per GPU "
                     "(126 MB L2)" % (frames_per_step * a.n_ant * N_SC * 8 / 2 ** 30)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def host_inputs(a, rank, torch, noise_scaler):
    """One step's worth of host-side inputs in PINNED memory, as a reference user would hand them over:
    receiver positions (float64), transmitted symbol values (uint8, = bit groups) and the additive
    noise on the data sub-carriers (complex64, unit variance scaled on the host like noise.py:64-66)."""
    F = a.frames
    rng = np.random.default_rng(2137 + rank)
    pos_t = torch.empty((F, 3), dtype=torch.float64, pin_memory=True)
    idx_t = torch.empty((F, N_SC), dtype=torch.uint8, pin_memory=True)
    nz_t = torch.empty((F, N_SC, 2), dtype=torch.float32, pin_memory=True)
    pos, idx, nz = pos_t.numpy(), idx_t.numpy(), nz_t.numpy().view(np.complex64).reshape(F, N_SC)
    pos[:, 0] = 212.0 + rng.uniform(-5, 5, F)
    pos[:, 1] = 212.0 + rng.uniform(-5, 5, F)
    pos[:, 2] = 1.5
    idx[:] = rng.integers(0, CONSTEL, (F, N_SC), dtype=np.uint8)
    # noise.py:64-66 with P = 42 * mean|sum_k alpha_k h_k v_k|^2 (mp_model.py:322,211); the scaler is the mean
    # over a few realisations measured on the device (per-frame values differ by <0.1 dB for this geometry)
    sigma = 0.5 * math.sqrt(2 * 42.0 * noise_scaler / 10 ** (snr_db_from_ebn0(a.ebn0) / 10))
    nz[:] = (rng.standard_normal((F, N_SC), dtype=np.float32)
             + 1j * rng.standard_normal((F, N_SC), dtype=np.float32)) * np.float32(sigma)
    keep = (pos_t, idx_t, nz_t)
    return dict(rx_pos=pos, tx_idx=idx, noise=nz), keep, pos.nbytes + idx.nbytes + nz.nbytes


def reference_measured():
    """The UNMODIFIED reference timed in the build container by tools/time_reference.py (it cannot travel to the GPU box):
    single-process Link.simulate and the mp.Process launcher pattern, frames/s with CPU model and core count."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_reference_timing.json")))
        return {"where": "%s, %s, %d vCPU (NOT this box)" % (d["where"], d["cpu"], d["vcpus"]),
                "single_process_mcnc_frames_per_s": d["single_process_mcnc"]["frames_per_s"],
                "single_process_cnc_frames_per_s": d["single_process_cnc"]["frames_per_s"],
                "multiprocess_mcnc_frames_per_s": d["multiprocess_mcnc"]["frames_per_s"],
                "multiprocess_processes": d["multiprocess_mcnc"]["processes"],
                "source": "profiles/r02_reference_timing.json (tools/time_reference.py)"}
    except Exception:
        return None


def secondary_workloads(eng, a, snr, peak):
    """The other fused shapes BASELINE.json names, measured in the same run after the headline legs (same GPU, same
    clocks): CNC receiver (configs[1]), two-user MU-MRT / ZF (configs[4]), SDR snapshots (configs[3]).  Every entry: frames
    (user-frames, snapshots) per second over the whole synchronous call incl. its D2H, the frame kernel's ms per launch
    and its fraction of the FP32 peak on the algorithmic FFT flops of that shape."""
    out = {}
    F, K, I = a.frames, a.n_ant, a.n_iter
    fft = 5.0 * N_FFT * math.log2(N_FFT)

    def run(name, fn, units, flops_per_unit, reps=3):
        fn()  # warm-up (scratch growth)
        eng.profile(True)
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        dt = time.perf_counter() - t0
        ms, n = eng.profile_read()
        eng.profile(False)
        k_ms = float(ms[2]) / max(int(n[2]), 1)
        out[name] = {"value": units * reps / dt, "unit": "frames/s", "units_per_call": units, "kernel_ms_per_launch": k_ms,
                     "frac": flops_per_unit * units / (k_ms * 1e-3) * 1e-12 / peak if k_ms > 0 else None}

    eng.configure(n_ant=K, channel="los", receiver="cnc", pa="softlim", ibo_db=0.0)
    run("cnc", lambda: eng.simulate(F, I, snr, seed=7), F, (2.0 * K + 2.0 * I) * fft)
    eng.configure(n_ant=K, channel="los", receiver="mcnc", pa="softlim", ibo_db=0.0)
    for tag in ("mrt", "zf"):
        run("mu_" + tag, lambda tag=tag: eng.mu_simulate(F // 2, I, snr, precoding=tag, seed=11), 2 * (F // 2),
            2.0 * K * (I + 1) * fft)
    run("sdr", lambda: eng.sdr(F, seed=3), F, 2.0 * K * fft)
    for v in list(out.values()):
        v["note"] = "device RNG, K=%d, I=%d, soft limiter IBO 0 dB" % (K, I)
    out["mu_mrt"]["unit"] = out["mu_zf"]["unit"] = "user-frames/s"
    out["sdr"]["unit"] = "snapshots/s"
    eng.configure(n_ant=K, channel="los", receiver="mcnc", pa="softlim", ibo_db=0.0)
    # the unfused building-block kernels (Transceiver.transmit + propagate stage by stage): HBM-bound, measured on 16384
    # device-resident rows (512 MiB per array, larger than L2) against the measured copy bandwidth
    hbm = float(measured_peaks().get("hbm_gbs", 6543.4))
    rows = 16384
    ms = eng.measure_stage_kernels(rows, 5)
    out["stages"] = {"bound": "hbm", "rows": rows, "row_samples": N_FFT, "peak_gbs": hbm,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy bandwidth)" if "hbm_gbs" in measured_peaks() else "fallback 6543.4",
                     "kernels": {k: {"ms_per_launch": ms[k], "algorithmic_bytes": b * rows,
                                     "achieved_gbs": b * rows / (ms[k] * 1e-3) * 1e-9, "frac": b * rows / (ms[k] * 1e-3) * 1e-9 / hbm}
                                 for k, b in zip(eng.STAGE_KERNELS, eng.STAGE_BYTES_PER_ROW)},
                     "note": "k_map_precode, k_fft4096_rows<+1>, k_pa<soft limiter>, k_fft4096_rows<-1>, k_propagate timed one by one "
                             "with CUDA events (mmimo_measure_stage_kernels); one stage call of the drop-in classes (64 rows, host "
                             "buffers) is latency-bound, this is the kernels' own rate; the peak is a measured COPY bandwidth (torch b.copy_(a)), not a hard ceiling: a streaming kernel with 16-byte accesses can exceed it by a few per cent (ncu DRAM peak of this GPU: 8.18 TB/s)"}
    return out


def run_gpu(a):
    import torch
    import torch.distributed as dist
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - this arm has no CPU fallback (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream(priority=-1)  # high priority: frame kernels before the library's background channel kernels
    torch.cuda.set_stream(stream)

    eng = LinkEngine(local)
    eng.configure(n_ant=a.n_ant, channel="los", receiver="mcnc", pa="softlim", ibo_db=0.0)
    eng.set_stream(stream.cuda_stream)
    snr = snr_db_from_ebn0(a.ebn0)
    I, F = a.n_iter, a.frames
    counters = torch.zeros(65, dtype=torch.int64, device="cuda")
    total = torch.zeros(65, dtype=torch.int64, device="cuda")

    def step(s, inject=None):
        """one batch of F frames on this GPU + the counter exchange (mp_model.py:221-222 -> one all-reduce)"""
        fid0 = (s * world + rank) * F
        eng.enqueue(F, I, snr, seed=2137, frame_id0=fid0, inject=inject)
        eng.counters_to_device(counters.data_ptr(), clear=True)
        if world > 1:
            dist.all_reduce(counters)
        total.add_(counters)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_result = torch.zeros(65, dtype=torch.int64, pin_memory=True)

    def timed(n_steps, first, inject=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for s in range(n_steps):
            step(first + s, inject)
            if inject is not None:
                # e2e: the result of EVERY step (the reduced error counts so far) is read back to pinned host memory,
                # stream-ordered and without blocking the enqueue of the next step - like a training loop logging its loss
                host_result.copy_(total, non_blocking=True)
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident arm ("value"): every input is generated on the GPU (Philox) ------------
    for s in range(a.warmup):
        step(s)
    barrier()
    total.zero_()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.launch_count
    eng.profile(True)
    ms_total = timed(a.steps, a.warmup)
    prof_ms, prof_n = eng.profile_read()
    eng.profile(False)
    launches = eng.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    tot = total.cpu().numpy()
    frames_total = a.steps * F * world
    value = frames_total / (ms_total * 1e-3)

    # ---- end-to-end arm ("e2e"): host buffers in, host counters out, every step ------------------
    probe = eng.simulate(8, 0, snr, seed=2137, dumps=("agc",))
    noise_scaler = float(np.mean(np.abs(probe["agc"].astype(np.complex128)) ** 2))
    inject, keep, h2d = host_inputs(a, rank, torch, noise_scaler)
    timed(1, 10_000, inject)  # warm-up of the injected path (staging buffers)
    total.zero_()
    del inject, keep
    e2e_steps = max(1, min(a.steps, 8))
    # ONE timed region over the e2e steps: every step copies its own inputs from pinned host memory (three cudaMemcpyAsync on
    # the library's copy stream, overlapping the previous step's frame kernel) and reads its result back (see timed()).
    # Each step gets its OWN host buffers, as a caller streaming fresh data would hand over.
    injects = [host_inputs(a, rank + 1000 * s, torch, noise_scaler) for s in range(min(e2e_steps, 3))]
    h2d = injects[0][2]
    e2e_frames = 0
    barrier()
    e2e_inj = [injects[s % len(injects)][0] for s in range(e2e_steps)]

    def e2e_region():
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for s in range(e2e_steps):
            step(20_000 + s, e2e_inj[s])
            host_result.copy_(total, non_blocking=True)
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    e2e_region()  # warm-up of the whole pipelined region (both staging slots)
    total.zero_()
    t_e2e = e2e_region()
    e2e_value = e2e_steps * F * world / (t_e2e * 1e-3)
    e2e_tot = total.cpu().numpy()

    if rank == 0:
        fl = flops_per_frame(a.n_ant, I)
        frame_ms, frame_n = float(prof_ms[2]), int(prof_n[2])
        achieved = fl * F * frame_n / (frame_ms * 1e-3) * 1e-12 if frame_ms > 0 else None
        peak = eng.measure_fp32_peak()
        nominal_peak = 148 * 128 * 2 * (measured_peaks().get("sm_max_mhz", 1965.0)) * 1e6 * 1e-12
        hbm_peak = measured_peaks().get("hbm_gbs")
        # algorithmic HBM bytes of the fused kernel per frame: the K channel rows re-read by every pass (the MRT
        # precoder is derived from them on the fly) + AGC / symbol-scale vectors
        bytes_frame = (I + 1) * a.n_ant * N_SC * 8 + (I + 1) * N_SC * 12
        # measured DRAM traffic of the kernel from the committed ncu capture, scaled to this launch size (same K, I only)
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "frame_kernel_traffic.json")))
            if tr["n_ant"] == a.n_ant and tr["n_iter"] == I:
                traffic = tr["dram_bytes_per_frame"] * F
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (complex64 compute, u64 error counters)", "data": "synthetic",
            "config": workload_config(a, F),
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 65 * 8,
                    "ber": (e2e_tot[:I + 1] / float(e2e_steps * F * world * 6 * N_SC)).tolist(),
                    "steps": e2e_steps,
                    "note": "one timed region of %d steps; every step: LinkEngine.enqueue(inject=pinned host rx_pos / tx symbols "
                            "/ noise, copied by the library's copy stream under the previous step's frame kernel) + counter "
                            "all-reduce + stream-ordered D2H of the 65 reduced counters into pinned host memory" % e2e_steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                         "traffic_note": "dram__bytes_read+write per launch from profiles/frame_kernel_traffic.json (ncu --set full, "
                                         "1184-frame launch) scaled to this launch's frame count; algorithmic HBM bytes per launch = %d" % (bytes_frame * F),
                         "kernel": {"r16": "k_link_frame<soft limiter>", "r64": "k_link_frame_r64<soft limiter>"}.get(
                             os.environ.get("MMIMO_FRAME_KERNEL", ""), "k_link_frame_t64<soft limiter>"), "launches_timed": frame_n,
                         "kernel_ms_per_launch": frame_ms / max(frame_n, 1),
                         "kernel_share_of_step": frame_ms / ms_total if ms_total else None,
                         "algorithmic_flops_per_frame": fl,
                         "peak_source": "FP32 FMA peak measured in this run on this GPU by mmimo_measure_fp32_peak "
                                        "(MEASURED_PEAKS.json carries no FP32 figure); nominal 148x128x2xf_max = %.1f"
                                        % nominal_peak,
                         "hbm": {"algorithmic_bytes_per_frame": bytes_frame,
                                 "achieved_gbs": bytes_frame * F * frame_n / (frame_ms * 1e-3) * 1e-9 if frame_ms > 0 else None,
                                 "peak_gbs": hbm_peak}},
            "kernel_ms": {"channel": float(prof_ms[0]), "prep": float(prof_ms[1]), "frame": frame_ms, "other": float(prof_ms[3])},
            "clocks": clocks,
            "ber": (tot[:I + 1] / float(frames_total * 6 * N_SC)).tolist(),
        }
        if world == 1 and not a.no_secondary:
            line["secondary"] = secondary_workloads(eng, a, snr, peak)
        if world == 1 and not a.no_cpu:
            line["cpu_baseline"] = cpu_baseline_single(a.n_ant, I, a.ebn0)
            line["cpu_baseline"]["reference_measured"] = reference_measured()
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # 4144 = 28 x 148: the frame kernel is persistent with one CTA per SM and one frame per CTA at a time, so a batch
    # that is a multiple of the SM count has no partially filled last wave (4096 would idle a third of the SMs for
    # 1/28 of the launch)
    ap.add_argument("--frames", type=int, default=4144, help="frames per step per GPU")
    ap.add_argument("--n-ant", type=int, default=64)
    ap.add_argument("--n-iter", type=int, default=5)
    ap.add_argument("--ebn0", type=float, default=15.0)
    ap.add_argument("--ref-frames-per-core", type=int, default=4)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary workloads (CNC, two-user, SDR)")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    if a.impl == "reference":
        run_reference(a)
    else:
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if world != a.gpus and world == 1 and a.gpus > 1:
            # convenience: re-launch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", "29571", os.path.abspath(__file__)] + sys.argv[1:]
            raise SystemExit(subprocess.call(cmd))
        run_gpu(a)


if __name__ == "__main__":
    main()
