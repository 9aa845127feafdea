"""Per-instruction stall summary of one kernel from an .ncu-rep (ncu --set full --import-source on)."""
import csv
import subprocess
import sys


def main(rep, pattern, ntop=24):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", ] + (["--kernel-name", f"regex:{pattern}"] if pattern else []),
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    i = hi[0]
    end = hi[1] - 1 if len(hi) > 1 else len(rows)
    h, body = rows[i], [r for r in rows[i + 1:end] if len(r) > 10]
    print(rows[i - 1][1][:100], len(body), "instructions")
    si, src = h.index("# Samples"), h.index("Source")
    stalls = [j for j, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[si] or 0) for r in body)
    agg = {h[j]: sum(int(r[j] or 0) for r in body) for j in stalls}
    print("samples", tot, sorted(agg.items(), key=lambda kv: -kv[1])[:8])
    for r in sorted(body, key=lambda r: -int(r[si] or 0))[:ntop]:
        st = sorted(((h[j], int(r[j] or 0)) for j in stalls if int(r[j] or 0) > 0), key=lambda kv: -kv[1])[:3]
        print(f"{r[si]:>6} {r[src][:64]:64s} {st}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 24)
