"""BASELINE configs[2]: element-order sweep ngl = 2..9 at ~16 M velocity dofs on the PER-CELL path (FreeSlip: element
matrices integrated for every cell, src/cases/base_problem.py:503-506).  One `bench.py --workload percell` line per
order: integration ms / TFLOP/s / flop per byte, K assembly ms and fraction of the HBM roofline -- the FLOP- vs
HBM-bound crossover.

    python tools/order_sweep.py [--fused 0|1] > profiles/r02_order_sweep_percell.jsonl
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
extra = sys.argv[1:]
for n in range(2, 10):
    nx = round(2828 / (n - 1))           # ~8.0 M nodes for every order
    subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "percell", "--nx", str(nx), "--ngl",
                    str(n), "--steps", "5", "--warmup", "3"] + extra, check=False)
