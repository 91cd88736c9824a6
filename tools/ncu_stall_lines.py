"""Where a kernel's warps wait, from the source page of an `ncu --set full --import-source on` report: the SASS
instructions with the most warp-stall samples and the reason that dominates each (development aid; the text it prints
is what profiles/*_ncu_summary.txt files carry under "stall samples by instruction").
    python tools/ncu_stall_lines.py report.ncu-rep [top_n]"""
import csv
import subprocess
import sys


def main(path, top_n=14):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: i for i, h in enumerate(hdr)}
    reasons = [h for h in hdr if h.startswith("stall_") and "(" not in h]
    recs = []
    for r in rows[hi + 1:]:
        if len(r) < len(hdr):
            continue
        try:
            n = int(r[col["# Samples"]] or 0)
        except ValueError:
            continue
        recs.append((n, r))
    total = sum(n for n, _ in recs) or 1
    by_reason = {h: sum(int(r[col[h]] or 0) for _, r in recs) for h in reasons}
    print(f"stall samples by reason (all {total} samples):",
          ", ".join(f"{h[6:]} {100.0 * v / total:.1f}%" for h, v in sorted(by_reason.items(), key=lambda x: -x[1])[:8]))
    print(f"stall samples by instruction (top {top_n} of {len(recs)} SASS instructions):")
    for n, r in sorted(recs, key=lambda x: -x[0])[:top_n]:
        top = max(reasons, key=lambda h: int(r[col[h]] or 0))
        src = " ".join(r[col["Source"]].split())
        print(f"  {100.0 * n / total:5.1f}%  {r[col['Address']][-5:]}  {src[:70]:70s} <- {top[6:]} {int(r[col[top]] or 0)}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 14)
