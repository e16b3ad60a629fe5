"""Attribute ncu SASS-level samples to CUDA source lines: python tools/ncu_lines.py <rep> <cubin> <kernel-substr> [top]"""
import csv, re, subprocess, sys
rep, cubin, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dem = sys.argv[5] if len(sys.argv) > 5 else kern          # substring of the demangled name in the ncu report
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# collect per-instruction line numbers of the kernel in order
lines, cur, infn = [], None, False
for l in dis:
    if l.startswith(".text."):
        infn = kern in l
        continue
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l): lines.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, curk = {}, None
for r in rows:
    if r and r[0] == "Kernel Name": curk = r[1]; blocks[curk] = []; continue
    if curk is not None: blocks[curk].append(r)
for k, rs in blocks.items():
    if dem not in k: continue
    hdr = rs[0]; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed")
    ins = rs[1:]
    print(k[:80], len(ins), "sass instrs; disasm has", len(lines))
    agg = {}
    for n, r in enumerate(ins):
        key = lines[n] if n < len(lines) else None
        a = agg.setdefault(key, [0, 0]); a[0] += int(r[si]); a[1] += int(r[ii])
    tot = sum(a[0] for a in agg.values())
    srcs = {}
    for key, (s, i) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        txt = ""
        if key:
            f = key[0]
            if f not in srcs:
                try: srcs[f] = open(f"task-aware-audio-coding-perceptual_b200/csrc/{f}").read().splitlines()
                except Exception: srcs[f] = []
            if key[1] - 1 < len(srcs[f]): txt = srcs[f][key[1] - 1].strip()
        print(f"{100*s/max(tot,1):5.1f}% {i:10d} {str(key):28s} {txt[:100]}")
    break
