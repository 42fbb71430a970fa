#!/usr/bin/env python
"""Per-source-line instruction counts of one captured kernel (ncu --import-source on):
  python tools/ncu_source_lines.py gpurun_out/x.ncu-rep [top]
Aggregates the 'Instructions Executed' column of the source page per (file, line)."""
import csv
import io
import subprocess
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fpath, hdr, agg = "", None, []
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            fpath = r[1].split("/")[-1]
        elif r and r[0] == "Line No":
            hdr = r
            ie, smp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        elif hdr and len(r) == len(hdr) and r[0] not in ("", "-"):
            try:
                agg.append((int(r[ie]), int(r[smp] or 0), fpath, r[0], r[1].strip()[:110]))
            except ValueError:
                pass
    tot = sum(a[0] for a in agg) or 1
    print(f"# {path}: {tot} warp instructions over {len(agg)} source lines")
    for n, s, f, ln, src in sorted(agg, reverse=True)[:top]:
        print(f"{n / tot:6.3f} {n:10d} smp={s:5d} {f}:{ln:>4s}  {src}")


if __name__ == "__main__":
    main()
