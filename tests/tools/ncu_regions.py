"""Dev tool: instruction / sample shares per source region of one `ncu --set full --import-source on` report.
usage: ncu_regions.py <report.ncu-rep> [top_lines]"""
import collections, csv, subprocess, sys

rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines())); h, u, r = rr[0], rr[1], rr[2]
for n in ("gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
          "launch__block_size", "smsp__thread_inst_executed_per_inst_executed.ratio"):
    if n in h: print(f"{n:60s} {r[h.index(n)]} {u[h.index(n)]}")
print("stall cycles per issue:", {n.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): round(float(r[i]), 2)
                                  for i, n in enumerate(h) if "issue_stalled" in n and n.endswith("per_issue_active.ratio") and float(r[i]) > 0.1})
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, hdr, agg = None, None, []
for row in csv.reader(src.splitlines()):
    if len(row) == 2 and row[0] in ("File Path", "File Name"): cur = row[1].split("/")[-1]
    elif len(row) > 8 and row[0] == "Line No": hdr = row; ii, si = row.index("Instructions Executed"), row.index("# Samples")
    elif hdr and len(row) > 8 and row[0].isdigit():
        try: agg.append((cur, int(row[0]), int(row[ii]), int(row[si]), row[1].strip()[:95]))
        except ValueError: pass
ti, ts = sum(a[2] for a in agg) or 1, sum(a[3] for a in agg) or 1
# regions of cabac_lockstep_enc.cuh by function
import os, re
here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "entropy_coding_b200", "csrc")
marks = {}
for f in ("cabac_lockstep_enc.cuh", "cabac_kernels.cu", "cabac_residual.cuh", "cabac_lockstep.cuh"):
    cur_fn, m = "(top)", []
    for ln, line in enumerate(open(os.path.join(here, f)), 1):
        g = re.match(r"^(?:template.*>\s*)?(?:__device__|__global__|struct|static)\b.*?(\w+)\s*(?:\(|\{|$)", line)
        if g and not line.startswith(" "): cur_fn = g.group(1)
        m.append(cur_fn)
    marks[f] = m
b, bs = collections.Counter(), collections.Counter()
for f, ln, i, s, t in agg:
    key = f + ":" + (marks[f][ln - 1] if f in marks and ln - 1 < len(marks[f]) else "")
    b[key] += i; bs[key] += s
print(f"total warp instructions {ti}, samples {ts}")
for k, v in b.most_common(22): print(f"  {k:55s} inst {100 * v / ti:5.1f}%  samples {100 * bs[k] / ts:5.1f}%")
for a in sorted(agg, key=lambda a: -a[2])[:top]: print(f"{100 * a[2] / ti:5.1f}% {100 * a[3] / ts:5.1f}%  {a[0]}:{a[1]} {a[4]}")
