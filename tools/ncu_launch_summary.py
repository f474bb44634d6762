"""Summarises an `ncu --metrics gpu__time_duration.sum[,dram__bytes_*] --csv` launch list per kernel name.
`--ours` keeps only this library's kernels (the capture window of `bench.py` starts with the synthetic-weight set-up, which
runs torch `at::` kernels outside the timed region)."""
import csv, sys, collections, re
ours = "--ours" in sys.argv
# --traffic-json=OUT: also write the per-launch DRAM traffic of the GEMM kernel family (bench.py's roofline.traffic reads it)
traffic_out = next((x.split("=", 1)[1] for x in sys.argv[1:] if x.startswith("--traffic-json=")), None)
src_csv = [x for x in sys.argv[1:] if not x.startswith("--")][0]
rows = list(csv.reader(l for l in open(src_csv) if l.startswith('"')))
hdr = rows[0]; H = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: dict(ids=set(), ns=0.0, rd=0.0, wr=0.0))
def val(r):
    v = float(r[H["Metric Value"]].replace(",", "")); u = r[H["Metric Unit"]]
    mult = {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9, "byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    return v * mult
skipped = set()
for r in rows[1:]:
    if ours and ("at::" in r[H["Kernel Name"]] or "cub::" in r[H["Kernel Name"]]):
        skipped.add(r[H["ID"]]); continue
    name = re.sub(r"\(.*", "", r[H["Kernel Name"]]).replace("b200sd::<unnamed>::", "").replace("void ", "")
    a = agg[name]; a["ids"].add(r[H["ID"]]); m = r[H["Metric Name"]]
    if m == "gpu__time_duration.sum": a["ns"] += val(r)
    elif m == "dram__bytes_read.sum": a["rd"] += val(r)
    elif m == "dram__bytes_write.sum": a["wr"] += val(r)
tot = sum(a["ns"] for a in agg.values())
print(f"# {len(rows)-1} metric rows, {sum(len(a['ids']) for a in agg.values())} launches, {tot/1e6:.3f} ms of kernel time (ncu per-launch times are cold-cache and serialised)")
if ours: print(f"# {len(skipped)} torch set-up launches (weight init / packing, outside the timed region) left out")
print(f"{'kernel':48s} {'launches':>8s} {'total ms':>9s} {'share':>6s} {'avg us':>8s} {'dram R MB/launch':>17s} {'dram W MB/launch':>17s}")
for name, a in sorted(agg.items(), key=lambda x: -x[1]["ns"]):
    n = len(a["ids"])
    print(f"{name[:48]:48s} {n:8d} {a['ns']/1e6:9.3f} {100*a['ns']/tot:5.1f}% {a['ns']/n/1e3:8.1f} {a['rd']/n/1e6:17.2f} {a['wr']/n/1e6:17.2f}")

if traffic_out:
    import json
    g = [a for n, a in agg.items() if n.startswith("gemm_tc_kernel")]
    n = sum(len(a["ids"]) for a in g)
    json.dump({"kernel": "gemm_tc_kernel", "launches_profiled": n, "dram_bytes_per_launch": sum(a["rd"] + a["wr"] for a in g) / max(n, 1),
               "source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none over "
                         "bench.py --steps 1 --warmup 3 (sd15_512_b8_ddim50_cfg); ncu flushes caches before every kernel, so this is a cold-cache figure",
               "csv": src_csv}, open(traffic_out, "w"), indent=1)
