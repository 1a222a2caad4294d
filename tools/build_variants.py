#!/usr/bin/env python
"""Build measurement variants of libmmimo_b200.so (same sources, different -D knobs) into variants/ and time them back
to back on ONE GPU box (the pool's B200s differ by ~3 % among themselves, so variants are never compared across boxes).

    python tools/build_variants.py build  name1:-DK=V,-DK2=V2  name2:...     # here (nvcc cross-compiles), in parallel
    python tools/build_variants.py time [--steps 10] [--env K=V ...] name1 name2 ...   # on the GPU box (gpurun)

`time` prints one line per (variant, env) with the frame-kernel ms per launch and frames/s from bench.py.
"""
import json
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VDIR = os.path.join(ROOT, "variants")


def build(specs):
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _build
    os.makedirs(VDIR, exist_ok=True)

    def one(spec):
        name, _, defs = spec.partition(":")
        defines = [d[2:] if d.startswith("-D") else d for d in defs.split(",") if d]
        out = os.path.join(VDIR, "lib_%s.so" % name)
        _build.build(force=True, defines=defines, out_path=out)
        return name, defines
    with ThreadPoolExecutor(max_workers=min(len(specs), os.cpu_count() or 1)) as ex:
        for name, defines in ex.map(one, specs):
            print("built", name, defines, flush=True)


def time_variants(argv):
    steps, envs, names = 10, [{}], []
    frames = None
    i = 0
    while i < len(argv):
        if argv[i] == "--steps":
            steps = int(argv[i + 1]); i += 2
        elif argv[i] == "--frames":
            frames = argv[i + 1]; i += 2
        elif argv[i] == "--env":  # --env K=a,b,c : cross product with what is there
            k, _, vs = argv[i + 1].partition("=")
            envs = [dict(e, **{k: v}) for e in envs for v in vs.split(",")]
            i += 2
        else:
            names.append(argv[i]); i += 1
    for rnd in range(2):
        for name in names:
            for e in envs:
                env = dict(os.environ, **e)
                if name != "default":
                    env["MMIMO_LIB"] = os.path.join(VDIR, "lib_%s.so" % name)
                cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--steps", str(steps), "--warmup", "3", "--no-cpu"]
                if frames:
                    cmd += ["--frames", frames]
                try:
                    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=75)
                except subprocess.TimeoutExpired:
                    print(name, e, 'TIMEOUT (hung kernel?)', flush=True)
                    continue
                try:
                    d = json.loads(res.stdout.strip().splitlines()[-1])
                    print("%-22s %-28s round %d  kernel %.3f ms  step %.3f ms  %.1f k frames/s  e2e %.1f k  frac %.4f  ber5 %.5f"
                          % (name, " ".join("%s=%s" % kv for kv in e.items()), rnd, d["roofline"]["kernel_ms_per_launch"],
                             d["ms_per_step"], d["value"] / 1e3, d["e2e"]["value"] / 1e3, d["roofline"]["frac"], d["ber"][-1]),
                          flush=True)
                except Exception:
                    print(name, e, "FAILED", res.stderr[-400:], flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2:])
    else:
        time_variants(sys.argv[2:])
