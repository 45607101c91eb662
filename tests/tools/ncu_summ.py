"""Dev tool: profiles/ summaries from one gpurun ncu capture.
usage: ncu_summ.py <tag> <launches.csv> <prof.ncu-rep> <bins_per_launch>"""
import collections, csv, json, subprocess, sys

tag, launches, rep, bins = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
rows = [r for r in csv.reader(l for l in open(launches) if not l.startswith("=="))]
hdr = rows[0]; ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    agg.setdefault(r[ki].split("(")[0], []).append(float(r[vi].replace(",", "")))
tot = sum(sum(v) for k, v in agg.items() if "cabac::" in k)
with open(f"profiles/{tag}_launches_summary.txt", "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none -c 400: python bench.py --no-cpu-baseline --no-e2e --steps 2 --warmup 3\n"
            "(cold-cache, serialised launches: compare shares, not absolutes; full list in %s_launches.csv)\n" % tag)
    for k, v in agg.items():
        share = 100 * sum(v) / tot if "cabac::" in k else 0
        f.write(f"{k[:70]:70s} n={len(v):3d} avg={sum(v) / len(v) / 1e6:9.4f} ms  share_of_cabac={share:5.1f}%\n")
open(f"profiles/{tag}_launches.csv", "w").write(open(launches).read())
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines())); h = rr[0]; ix = {n: i for i, n in enumerate(h)}
facts = {}
for r in rr[2:]:
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("cabac::", "").split("<")[0].strip()
    g = lambda m: float(r[ix[m]].replace(",", ""))
    unit = lambda m: rr[1][ix[m]]
    def by(m):
        v, u = g(m), unit(m)
        return int(v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u])
    inst = g("smsp__inst_executed.sum")
    facts[name] = {
        "warp_instr_per_bin": round(inst / bins, 2), "warp_instr_per_launch": inst, "bins_per_launch": bins,
        "dram_bytes_per_launch": by("dram__bytes_read.sum") + by("dram__bytes_write.sum"),
        "dram_read_bytes": by("dram__bytes_read.sum"), "dram_write_bytes": by("dram__bytes_write.sum"),
        "issue_active_pct": g("smsp__issue_active.avg.pct") if "smsp__issue_active.avg.pct" in ix else None,
        "registers_per_thread": int(g("launch__registers_per_thread")),
        "achieved_occupancy_pct": g("sm__warps_active.avg.pct_of_peak_sustained_active"),
        "shared_bank_conflicts": int(g("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")),
        "thread_inst_per_warp_inst": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "duration_ms_under_ncu": g("gpu__time_duration.sum") / (1e6 if unit("gpu__time_duration.sum") in ("ns", "nsecond") else 1),
        "stall_cycles_per_issue": {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): round(float(r[i]), 2)
                                   for i, k in enumerate(h) if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("per_issue_active.ratio") and float(r[i]) >= 0.05},
        "source": f"profiles/{tag}_ncu_details.txt (ncu --set full --clock-control none, 64-frame 1080p batch, 8640 substreams)"}
det = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
open(f"profiles/{tag}_ncu_details.txt", "w").write("\n".join(l for l in det.splitlines() if l.strip()))
cur = None
for line in det.splitlines():  # "Issue Slots Busy" lives on the details page only
    for name in facts:
        if name + "(" in line or name + "<" in line:
            cur = name
    if cur and "Issue Slots Busy" in line:
        facts[cur]["issue_active_pct"] = float(line.split()[-1])
json.dump(facts, open("profiles/kernel_facts.json", "w"), indent=1)
print(json.dumps(facts, indent=1))
