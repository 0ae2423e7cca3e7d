"""A/B of the CTA-pair convolution kernel (mode 2: 256-filter tiles, 3: 128-filter tiles, 1: automatic) against the
single-CTA one (0) on the 256/512-filter shapes (development tool)."""
import subprocess, sys
for on in sys.argv[1:] or ("1", "2", "3", "0"):
    print("pair mode", on, flush=True)
    subprocess.run([sys.executable, "-c",
                    "import sys; sys.path.insert(0,'.'); sys.argv=['x','4096','conv']\n"
                    "from iit_b200 import _lib\n"
                    f"_lib.lib.iit_conv_pair_enable({on})\n"
                    "exec(open('scratch/bench_kernels.py').read())"], check=False)
