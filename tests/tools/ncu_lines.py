"""Dev tool: top source lines of an `ncu --page source --csv --print-source cuda,sass` dump."""
import csv
import sys


def main(path, top=45):
    rows = list(csv.reader(open(path, errors="replace")))
    cur, hdr, agg = None, None, []
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif len(r) > 8 and r[0] == "Line No":
            hdr = r
            ii, si = r.index("Instructions Executed"), r.index("# Samples")
        elif hdr and len(r) > 8 and r[0].isdigit():
            try:
                agg.append((int(r[ii]), int(r[si]), cur, int(r[0]), r[1].strip()[:100]))
            except ValueError:
                pass
    tot_i = sum(a[0] for a in agg) or 1
    tot_s = sum(a[1] for a in agg) or 1
    print(f"total warp instructions {tot_i}, samples {tot_s}")
    for a in sorted(agg, reverse=True)[:top]:
        print(f"{100 * a[0] / tot_i:5.1f}% inst {100 * a[1] / tot_s:5.1f}% smp  {a[2]}:{a[3]:<4} {a[4]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)
