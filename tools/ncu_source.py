"""Per-source-line instruction breakdown of an .ncu-rep (needs --import-source on / -lineinfo). Usage: ncu_source.py rep [topN]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur = None; hdr = None; out = []
for r in rows:
    if len(r) >= 2 and r[0] in ("File Path", "File Name"): cur = r[1].split('/')[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        try: ie = int(d["Instructions Executed"]); te = int(d["Thread Instructions Executed"]); smp = int(d["# Samples"])
        except Exception: continue
        out.append((cur, int(r[0]), r[1].strip()[:84], ie, te, smp))
tot = sum(o[3] for o in out); tots = max(1, sum(o[5] for o in out))
print(f"total warp instructions {tot:.4g}, stall samples {tots}")
byfile = {}
for o in out:
    a = byfile.setdefault(o[0], [0, 0, 0]); a[0] += o[3]; a[1] += o[4]; a[2] += o[5]
for f, a in byfile.items():
    if a[0]: print(f"  {f:28s} inst {100*a[0]/tot:5.1f}%  threads/inst {a[1]/max(a[0],1):5.1f}  samples {100*a[2]/tots:5.1f}%")
out.sort(key=lambda o: -o[3])
for o in out[:top]:
    print(f"{o[0][:20]:20s} {o[1]:4d} inst {100*o[3]/tot:5.1f}% thr/inst {o[4]/max(o[3],1):5.1f} smp {100*o[5]/tots:5.1f}%  {o[2]}")
