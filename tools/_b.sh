python bench.py --no-cpu --steps 10 > gpurun_out/bb.log 2>&1; python - <<EOF
import json
d=json.loads(open("gpurun_out/bb.log").read().strip().splitlines()[-1])
print("RESULT", round(d["value"]), d["ms_per_step"], {k: round(v/13,3) for k,v in d["kernel_ms"].items()})
EOF
