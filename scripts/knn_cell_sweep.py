"""Sweep the candidates-per-cell parameter of the grid K-NN (MACBF_KNN_CELL) on configs[1]."""
import os, subprocess, sys
for p in (2.0, 3.0, 4.0, 5.0, 6.0, 7.0, 9.0):
    env = dict(os.environ, MACBF_KNN_CELL=str(p))
    out = subprocess.run([sys.executable, "scripts/kernel_times.py"], env=env, capture_output=True, text=True).stdout
    line = [l for l in out.splitlines() if l.startswith("neighbor_topk")]
    print(p, line[0] if line else out[-300:])
