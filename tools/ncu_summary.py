"""Summarise an .ncu-rep (CPU side): python tools/ncu_summary.py <rep> [kernel-substr]  -> key metrics + opcode mix + stall mix."""
import collections, csv, re, subprocess, sys

rep = sys.argv[1]
sub = sys.argv[2] if len(sys.argv) > 2 else ""
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.sum", "sm__inst_executed_pipe_tensor.sum",
        "sm__inst_executed_pipe_uniform.sum", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    if sub and sub not in r[ki]:
        continue
    print("==", r[ki][:100])
    d = dict(zip(hdr, r))
    u = dict(zip(hdr, units))
    for k in KEYS:
        if k in d:
            print(f"  {k:75s} {d[k]:>16s} {u[k]}")
    st = {h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]: float(d[h]) for h in hdr
          if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and d[h]}
    print("  stalls (warps per issue):", ", ".join(f"{k} {v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:10]))
    break
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for r in csv.reader(src.splitlines()):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if cur is not None:
        cur["rows"].append(r)
for b in blocks:
    if sub and sub not in b["name"]:
        continue
    h = b["rows"][0]
    si, ie, ss = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    op, smp = collections.Counter(), collections.Counter()
    for r in b["rows"][1:]:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[si])
        o = m.group(2).split(".")[0] if m else "?"
        op[o] += int(r[ie]); smp[o] += int(r[ss])
    tot, ts = sum(op.values()), max(sum(smp.values()), 1)
    print(f"  SASS: {len(b['rows']) - 1} static instrs, {tot} executed")
    print("  opcode mix:", ", ".join(f"{o} {100 * c / tot:.1f}%" for o, c in op.most_common(16)))
    break
