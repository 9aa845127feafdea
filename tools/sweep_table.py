"""profiles/r02_order_sweep_percell.md from the order-sweep .jsonl files: tools/order_sweep.py --fused 1, --fused 0 and
(optional) --fused 1 under PYN_AFF_NO_SIG=1 (the slot-flat kernel)."""
import json
import sys

fused, mat, flat = {}, {}, {}
for path, dst in ((sys.argv[1], fused), (sys.argv[2], mat), (sys.argv[3] if len(sys.argv) > 3 else None, flat)):
    if path is None:
        continue
    for l in open(path):
        if l.startswith("{"):
            d = json.loads(l)
            dst[int(d["config"]["workload"].split("ngl ")[1].split(" ")[0])] = d
print("# BASELINE configs[2]: element-order sweep on the per-cell path, ~16 M velocity dofs, one B200\n")
print("`python tools/order_sweep.py --fused 1|0` = `bench.py --workload percell --nx round(2828/(ngl-1)) --ngl n` per order "
      "(FreeSlip.buildKLEMats: element matrices of EVERY cell + K, Krhs, Rw values on the prebuilt pattern; CUDA events, "
      "5 timed steps).  `frac` = (8 nnz(K) + 8 (rows+1) + 64 cells) B / time to produce K from corner coordinates / "
      "6 543 GB/s.\n")
print("## materialise + gather (`-mat_fused 0`: general straight-sided quads)\n")
print("| ngl | cells | nnz(K) | step ms | `elem_kle_kernel` ms | TFLOP/s (sum-factorised flops) | of 37 TF FP64 | GB/s written | flop/B | "
      "K gather ms | K frac of HBM | G DOF/s |\n|---|---|---|---|---|---|---|---|---|---|---|---|")
for n in sorted(mat):
    d = mat[n]; i = d["integration"]; p = d["phases_ms"]; c = d["config"]
    print(f"| {n} | {c['cells']:,} | {c['nnz']['K']:,} | {d['ms_per_step']:.2f} | {p['elem']:.2f} | {i['tflops']:.2f} | "
          f"{i['fp64_frac']:.3f} | {i['gbs']:.0f} | {i['flop_per_byte']:.1f} | {p['K']:.2f} | {d['roofline']['frac']:.2f} | "
          f"{d['value'] / 1e9:.2f} |")
def fused_table(rows):
    print("| ngl | step ms | coefficients ms | K ms | K frac of HBM | Rw ms | G DOF/s | speed-up of the step |\n|---|---|---|---|---|---|---|---|")
    for n in sorted(rows):
        d = rows[n]; p = d["phases_ms"]
        print(f"| {n} | {d['ms_per_step']:.2f} | {p['elem']:.2f} | {p['K']:.2f} | {d['roofline']['frac']:.2f} | {p['Rw']:.2f} | "
              f"{d['value'] / 1e9:.2f} | {mat[n]['ms_per_step'] / d['ms_per_step']:.1f}x |")


print("\n## fused integrate-and-assemble, signature-tile kernel (default on structured meshes of parallelogram cells)\n")
fused_table(fused)
if flat:
    print("\n## fused integrate-and-assemble, slot-flat kernel (`PYN_AFF_NO_SIG=1`; meshes without row signatures)\n")
    fused_table(flat)
print("""
Reading.  The arithmetic intensity of the sum-factorised integration (its flops / the `64 n^4` B of element matrices it
writes) crosses the machine balance of the B200 (37 TFLOP/s FP64 / 6.5 TB/s = 5.7 flop/B) at **ngl 4**: below it an
ideal kernel would be HBM-bound, from ngl 5 on FLOP-bound.  The kernel reaches neither roof: at most 6.2 TFLOP/s (0.17
of the FP64 pipe) and 0.08 of the HBM bandwidth; ncu (`r02_ncu_elem_kle.md`) shows why -- at ngl 5 a third of the warp
stalls are CTA barriers (eight per element) and a fifth shared-memory operand latency (every multiply-add takes two
operands from shared memory), at ngl 2 instruction fetch (a long kernel run by one warp per 8x8 matrix).  On general
quads the integration therefore dominates the per-cell step 5-10x at every order.  The closed form on parallelogram
cells removes the contraction and the element matrices altogether: the step becomes the CSR store stream plus the
evaluation of the entries from 13 coefficients and a 64-byte product record per 2x2 block.  The slot-flat kernel pays
~350 instructions per 32 blocks for look-ups and is issue-bound at 0.1-0.3 of the HBM roofline; the signature-tile
kernels (`r02_ncu_fused_assembly.md`) bring that to 80-120 and reach **0.48-0.75** for ngl >= 4: 0.51 / 0.68 / 0.75 at
ngl 4 / 6 / 8, whose rows are whole 32-byte sectors or 128-byte lines, and 0.48 / 0.58 / 0.57 at ngl 5 / 7 / 9, where
every second row starts mid-sector and the line-aligned variant of the kernel is used (0.44 / 0.51 / 0.49 without
it); lowest for the short rows of ngl 2-3 (0.19 / 0.30; 9-25 column nodes: the per-tile set-up is not amortised).  Its cost per
stored entry is flat in ngl, so this path has no crossover: it is a store
stream at every order.""")
