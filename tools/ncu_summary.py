"""Markdown summary of the kernels in an .ncu-rep (ncu --set full --import-source on): headline metrics, warp-stall
breakdown and the hottest SASS instructions.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep [kernel-substring] > profiles/x.md
"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_fp64.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]


def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout


def main(rep, sub=""):
    raw = list(csv.reader(run([rep, "--page", "raw", "--csv"]).splitlines()))
    h, units = raw[0], raw[1]
    print(f"# ncu summary of `{rep.split('/')[-1]}`\n")
    for r in raw[2:]:
        d = dict(zip(h, r))
        if sub and sub not in d.get("Kernel Name", ""):
            continue
        print(f"## `{d['Kernel Name'][:110]}`\n\n| metric | value | unit |\n|---|---|---|")
        for k in KEYS:
            if k in d and d[k] != "":
                print(f"| {k} | {d[k]} | {units[h.index(k)]} |")
        print()
    src = list(csv.reader(run([rep, "--page", "source", "--csv"]).splitlines()))
    starts = [i for i, r in enumerate(src) if r and r[0] == "Kernel Name"]
    seen = set()
    for n, i in enumerate(starts):
        name = src[i][1]
        if (sub and sub not in name) or name in seen:
            continue
        seen.add(name)
        end = starts[n + 1] if n + 1 < len(starts) else len(src)
        hh, body = src[i + 1], [r for r in src[i + 2:end] if len(r) > 10]
        si, so, ie = hh.index("# Samples"), hh.index("Source"), hh.index("Instructions Executed")
        stalls = [j for j, c in enumerate(hh) if c.startswith("stall_") and "Not Issued" not in c]
        tot = sum(int(r[si] or 0) for r in body) or 1
        agg = sorted(((hh[j], sum(int(r[j] or 0) for r in body)) for j in stalls), key=lambda kv: -kv[1])
        print(f"### warp stalls, `{name[:90]}`\n\n{len(body)} SASS instructions, "
              f"{sum(int(r[ie] or 0) for r in body):,} warp instructions executed, {tot:,} samples\n")
        print("| stall | share |\n|---|---|")
        for k, v in agg[:7]:
            print(f"| {k} | {100 * v / tot:.1f} % |")
        print("\n| samples | executed | instruction | top stalls |\n|---|---|---|---|")
        for r in sorted(body, key=lambda r: -int(r[si] or 0))[:12]:
            st = sorted(((hh[j], int(r[j] or 0)) for j in stalls if int(r[j] or 0) > 0), key=lambda kv: -kv[1])[:2]
            print(f"| {r[si]} | {r[ie]} | `{r[so].strip()[:70]}` | {', '.join(f'{a} {b}' for a, b in st)} |")
        print()


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
