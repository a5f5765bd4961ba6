"""scratch: each GPU test file in its own process with a time limit (a hang must not eat the GPU budget)"""
import subprocess, sys
for args, limit in ((["tests/test_gpu_arith.py"], 200), (["tests/test_gpu_parity.py", "-k", "cooperative"], 200)):
    try:
        r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu"] + args, capture_output=True, text=True, timeout=limit)
        print(args, "rc", r.returncode, r.stdout.strip().splitlines()[-6:], flush=True)
    except subprocess.TimeoutExpired:
        print(args, "TIMEOUT", flush=True)
