"""profiles/traffic.json from an ncu metrics pass of the bench command (CPU side):

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum \\
        --clock-control none --csv --log-file gpurun_out/traffic_<tag>.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra
    python tools/make_traffic.py gpurun_out/traffic_<tag>.csv profiles/traffic_<tag>.csv

Sums, over the kernels of each leg in the LAST step of the capture, the DRAM bytes and the warp instructions per launch."""
import collections, csv, json, os, shutil, sys
src, keep = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(src)) if len(r) > 10]
h = rows[0]
ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
launches = collections.OrderedDict()
for r in rows[1:]:
    launches.setdefault(int(r[ii]), {"name": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
LEGS = {"loss_fused": ("mrstft_stream_kernel", "mrstft_stream_small_kernel", "mrstft_combine_kernel", "mrstft_stream_sums_kernel"),
        "frontend": ("stft_frontend_kernel",)}
ids = sorted(launches)
out = {}
for leg, kernels in LEGS.items():
    tot = {"traffic": 0.0, "inst": 0.0, "us": 0.0, "kernels": {}}
    for kname in kernels:
        last = [i for i in ids if kname + "<" in launches[i]["name"] or kname + "(" in launches[i]["name"]]
        if not last:
            continue
        L = launches[last[-1]]
        b = L.get("dram__bytes_read.sum", 0) + L.get("dram__bytes_write.sum", 0)
        tot["traffic"] += b; tot["inst"] += L.get("smsp__inst_executed.sum", 0); tot["us"] += L.get("gpu__time_duration.sum", 0) / 1e3
        tot["kernels"][kname] = {"dram_bytes": int(b), "inst": int(L.get("smsp__inst_executed.sum", 0)), "us": L.get("gpu__time_duration.sum", 0) / 1e3}
    out[leg] = {"traffic": int(tot["traffic"]), "inst": int(tot["inst"]), "kernels": tot["kernels"],
                "source": f"{keep} (ncu metrics pass of `bench.py --steps 2 --warmup 3`, last step; cold-cache, serialised launches)"}
shutil.copyfile(src, keep)
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
json.dump(out, open(os.path.join(root, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
