"""Annotated per-line listing of an .ncu-rep: warp instructions, lanes and stall samples beside the local source text.
Usage: ncu_annotate.py rep [min_percent] [git_rev]   (source text read from the working tree, or from `git show rev:`)"""
import csv, io, subprocess, sys, glob, os
rep = sys.argv[1]; minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.15; rev = sys.argv[3] if len(sys.argv) > 3 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur = None; hdr = None; out = {}
for r in rows:
    if len(r) >= 2 and r[0] in ("File Path", "File Name"): cur = r[1].split('/')[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        try: ie = int(d["Instructions Executed"]); te = int(d["Thread Instructions Executed"]); smp = int(d["# Samples"])
        except Exception: continue
        a = out.setdefault(cur, {}).setdefault(int(r[0]), [0, 0, 0]); a[0] += ie; a[1] += te; a[2] += smp
tot = sum(a[0] for f in out.values() for a in f.values()); tots = max(1, sum(a[2] for f in out.values() for a in f.values()))
def text(fname):
    hits = [p for p in glob.glob("volume_path_tracer_b200/csrc/**/" + fname, recursive=True)]
    if not hits: return []
    if rev: return subprocess.run(["git", "show", f"{rev}:{hits[0]}"], capture_output=True, text=True).stdout.splitlines()
    return open(hits[0]).read().splitlines()
print(f"total warp instructions {tot:.4g}, samples {tots}")
for f, lines in out.items():
    src = text(f); ft = sum(a[0] for a in lines.values())
    if ft < tot * 0.002: continue
    print(f"==== {f}: {100*ft/tot:.1f}% of warp instructions")
    for ln in sorted(lines):
        a = lines[ln]
        if a[0] < tot * minp / 100: continue
        t = src[ln - 1].strip()[:110] if 0 < ln <= len(src) else "?"
        print(f"{ln:4d} {100*a[0]/tot:5.2f}% lanes {a[1]/max(a[0],1):4.1f} smp {100*a[2]/tots:5.2f}% | {t}")
