"""Per-instruction table of the global loads of an .ncu-rep (SASS page): warp instructions, threads, L1 tag requests, theoretical L2 sectors.
Usage: python tools/ncu_loads.py rep [min_share_percent]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(r for r in rows if r and r[0] == "Address")
col = {h: i for i, h in enumerate(hdr)}
loads = []
tot_inst = 0
for r in rows:
    if len(r) != len(hdr) or r[0] == "Address": continue
    try: ie = int(r[col["Instructions Executed"]])
    except ValueError: continue
    tot_inst += ie
    s = r[col["Source"]].strip()
    if s.startswith("LDG") or s.startswith("@") and " LDG" in s or "ATOM" in s or "RED" in s or s.startswith("STG"):
        loads.append((s, ie, int(r[col["Thread Instructions Executed"]]), int(r[col["L1 Tag Requests Global"]]), int(r[col["L2 Theoretical Sectors Global"]]), int(r[col["L2 Theoretical Sectors Global Ideal"]]), int(r[col["# Samples"]])))
tl = sum(l[1] for l in loads); ts = sum(l[4] for l in loads)
print(f"total warp instructions {tot_inst:.4g}; global memory instructions {tl:.4g} ({100*tl/tot_inst:.1f} %), theoretical L2 sectors {ts:.4g}")
print(f"{'instruction':70s} {'warp-inst':>10s} {'lanes':>6s} {'tags/inst':>9s} {'sect/inst':>9s} {'sect/thread':>11s} {'ideal/inst':>10s} {'share of sectors':>8s}")
for s, ie, te, tags, sec, ideal, smp in loads:
    if ie == 0 or sec < ts * minp / 100: continue
    print(f"{s[:70]:70s} {ie:10.3g} {te/ie:6.2f} {tags/ie:9.2f} {sec/ie:9.2f} {sec/te:11.3f} {ideal/ie:10.2f} {100*sec/ts:7.1f}%")
