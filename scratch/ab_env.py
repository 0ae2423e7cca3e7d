"""A/B of an environment switch over alternating bench runs: python scratch/ab_env.py NAME v1 v2 (development tool)."""
import json, os, subprocess, sys
name, vals = sys.argv[1], sys.argv[2:]
for val in vals * 2:
    env = dict(os.environ, **{name: val})
    out = subprocess.run([sys.executable, "bench.py", "--steps", "10", "--warmup", "3"], capture_output=True, text=True, env=env)
    d = json.loads(out.stdout.strip().splitlines()[-1])
    print(name, val, round(d["ms_per_step"], 3), round(d["e2e"]["value"]), flush=True)
