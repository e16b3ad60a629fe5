"""Per-region stall profile of one kernel from an .ncu-rep's source page (CPU side):
    python tools/ncu_hot.py <rep> [region_size]
prints, for consecutive regions of `region_size` SASS instructions: executed warp-instructions, stall samples and the
three largest stall reasons, plus the dominant opcodes -- enough to tell the phases of an unrolled kernel apart."""
import collections, csv, re, subprocess, sys
rep = sys.argv[1]
reg = int(sys.argv[2]) if len(sys.argv) > 2 else 500
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
h = rows[1]
si, ie, ss = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
stall = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
body = rows[2:]
tot_s = sum(int(r[ss]) for r in body); tot_e = sum(int(r[ie]) for r in body)
print(f"total executed {tot_e}, samples {tot_s}")
for a in range(0, len(body), reg):
    blk = body[a:a + reg]
    e = sum(int(r[ie]) for r in blk); s = sum(int(r[ss]) for r in blk)
    if e == 0 and s == 0:
        continue
    st = collections.Counter()
    op = collections.Counter()
    for r in blk:
        for i, c in stall:
            st[c[6:]] += int(r[i] or 0)
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[si])
        if m: op[m.group(2)] += int(r[ie])
    top = ", ".join(f"{k} {100 * v / max(s, 1):.0f}%" for k, v in st.most_common(4))
    ops = ", ".join(f"{k} {100 * v / max(e, 1):.0f}%" for k, v in op.most_common(5))
    print(f"[{a:6d}] exec {100 * e / tot_e:5.1f}%  samples {100 * s / tot_s:5.1f}%  ratio {(s / tot_s) / max(e / tot_e, 1e-9):4.2f} | {top} | {ops}")
