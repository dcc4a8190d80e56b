#!/usr/bin/env python3
"""ncu_source.py REPORT.ncu-rep [kernel-substring] -- per-opcode summary of an `ncu --set full --import-source on` report:
instructions executed, shared-memory wavefronts (actual / ideal / excess) and stall samples per SASS opcode, and the
instructions with the most excess wavefronts.  Reads the report with `ncu -i ... --page source --csv` (no GPU needed).
Used for the hot-line tables in DESIGN.md and profiles/*_sass.txt."""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    lines = out.splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
    rows = list(csv.DictReader(lines[start:]))
    ops = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
    tot = [0, 0, 0, 0, 0]
    for r in rows:
        try:
            inst = int(r["Instructions Executed"])
        except (ValueError, KeyError):
            continue
        src = r["Source"].strip()
        toks = src.split()
        op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
        op = ".".join(op.split(".")[:3])
        wf, ideal = int(r["L1 Wavefronts Shared"] or 0), int(r["L1 Wavefronts Shared Ideal"] or 0)
        smp = int(r["# Samples"] or 0)
        for acc in (ops[op], tot):
            acc[0] += inst; acc[1] += wf; acc[2] += ideal; acc[3] += smp; acc[4] += 1
    print(f"{'opcode':28s} {'static':>6s} {'warp-inst':>12s} {'%':>6s} {'smem wavefronts':>16s} {'ideal':>12s} {'samples%':>8s}")
    for op, v in sorted(ops.items(), key=lambda kv: -kv[1][0])[:40]:
        print(f"{op:28s} {v[4]:6d} {v[0]:12d} {100*v[0]/tot[0]:6.2f} {v[1]:16d} {v[2]:12d} {100*v[3]/max(tot[3],1):8.2f}")
    print(f"{'TOTAL':28s} {tot[4]:6d} {tot[0]:12d} {100.0:6.2f} {tot[1]:16d} {tot[2]:12d}")
    print("\nmost excess shared-memory wavefronts:")
    ex = sorted(rows, key=lambda r: -(int(r.get("L1 Wavefronts Shared Excessive") or 0)))[:12]
    for r in ex:
        print(f"  {r['Source'].strip()[:70]:70s} inst {r['Instructions Executed']:>10s} wf {r['L1 Wavefronts Shared']:>10s} excess {r['L1 Wavefronts Shared Excessive']:>10s}")


if __name__ == "__main__" and not (len(sys.argv) > 2 and sys.argv[2] == "lines"):
    main()


def by_line(rep, top=40):
    """Per CUDA-C source line: warp instructions, stall samples, shared-memory wavefronts (ncu ... --print-source cuda,sass)."""
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[hi]
    nm = len(hdr) - 4                                   # metric columns behind (line, source, address, sass)
    names = hdr[4:]
    ii, si, wi = names.index("Instructions Executed"), names.index("# Samples"), names.index("L1 Wavefronts Shared")
    lines = []
    for r in rows[hi + 1:]:
        if r and r[0].isdigit() and len(r) >= nm + 2:   # a source line (its text may hold commas: count from the right)
            m = r[-nm:]
            try:
                lines.append((int(r[0]), ",".join(r[1:len(r) - nm]).strip()[:90], int(m[ii] or 0), int(m[si] or 0), int(m[wi] or 0)))
            except ValueError:
                pass
    tot = sum(x[2] for x in lines) or 1
    tots = sum(x[3] for x in lines) or 1
    print(f"{'line':>5s} {'warp-inst':>12s} {'%':>6s} {'samples%':>8s} {'smem wf':>11s}  source")
    for ln, src, inst, smp, wf in sorted(lines, key=lambda x: -x[2])[:top]:
        print(f"{ln:5d} {inst:12d} {100*inst/tot:6.2f} {100*smp/tots:8.2f} {wf:11d}  {src}")


if __name__ == "__main__" and len(sys.argv) > 2 and sys.argv[2] == "lines":
    by_line(sys.argv[1])
