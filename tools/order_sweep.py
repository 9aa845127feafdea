"""BASELINE configs[2]: element-order sweep ngl = 2..9 at ~8.0 M nodes (16 M velocity dofs), one B200.
Runs bench.py per order (assembly step + CG legs) and prints one table row per order; raw lines -> the given file."""
import json
import subprocess
import sys

out_path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/order_sweep.jsonl"
rows = []
with open(out_path, "w") as f:
    for n in range(2, 10):
        nx = round(2828 / (n - 1))
        cmd = [sys.executable, "bench.py", "--nx", str(nx), "--ngl", str(n), "--steps", "3", "--warmup", "3",
               "--no-cpu-baseline", "--no-e2e", "--ksp-its", "20"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
        if r.returncode != 0 or not line.startswith("{"):
            print(f"ngl {n} nx {nx}: FAILED rc={r.returncode}\n{r.stderr[-1500:]}")
            continue
        f.write(line + "\n")
        d = json.loads(line)
        k = d["ksp"]
        e = k.get("element_operator")
        rows.append((n, nx, d["config"]["dof_rows"], d["config"]["nnz_rank0"]["K"], d["ms_per_step"], d["value"] / 1e9,
                     d["roofline"]["frac"], k["csr_stream"]["ms_per_iteration"], k["csr_stream"]["spmv"]["frac"],
                     e["ms_per_iteration"] if e else float("nan"), e["spmv"]["ms"] if e else float("nan")))
print("| ngl | nx | dof rows | nnz(K) | asm ms/step | G DOF/s | K-asm frac of HBM | CG it ms (CSR) | SpMV frac | CG it ms (elem. op.) | elem. op. product ms |")
print("|---|---|---|---|---|---|---|---|---|---|---|")
for r in rows:
    print("| %d | %d | %d | %d | %.2f | %.2f | %.2f | %.2f | %.2f | %.2f | %.2f |" % r)
