"""Summarise an ncu source page CSV (ncu -i rep --page source --csv): top instructions by stall samples with the
dominant stall reason, plus per-source-line totals when SASS->source mapping is present."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
S = hdr.index("# Samples"); SRC = hdr.index("Source")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = 0; out = []
reason_tot = collections.Counter()
for r in rows[hi + 1:]:
    if len(r) <= S: continue
    try: n = int(r[S])
    except ValueError: continue
    tot += n
    st = {hdr[i]: int(r[i] or 0) for i in stall_cols if r[i] not in ("", "0")}
    for k, v in st.items(): reason_tot[k] += v
    out.append((n, r[SRC], st))
print("total samples", tot)
print("by reason:", [(k, f"{100*v/max(tot,1):.1f}%") for k, v in reason_tot.most_common(8)])
for n, src, st in sorted(out, key=lambda x: -x[0])[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    top = sorted(st.items(), key=lambda x: -x[1])[:2]
    print(f"{100*n/max(tot,1):5.1f}%  {src[:90]:90s} {top}")
