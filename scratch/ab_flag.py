"""A/B of an ops-level switch inside one process-per-setting bench run: python scratch/ab_flag.py BN_RELU_MASK (development tool)."""
import json, subprocess, sys
flag = sys.argv[1]
for val in ("True", "False", "True", "False"):
    code = (f"import sys; sys.argv=['bench.py','--steps','10','--warmup','3']; sys.path.insert(0,'.')\n"
            f"from iit_b200 import ops; ops.{flag} = {val}\n"
            "import runpy; runpy.run_path('bench.py', run_name='__main__')")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True).stdout.strip().splitlines()
    d = json.loads(out[-1])
    print(flag, val, round(d["ms_per_step"], 3), round(d["e2e"]["value"]), flush=True)
