"""Instruction-cache view of an .ncu-rep (SASS page): how many of the kernel's static instructions are executed, and how the executed ones cluster.
Usage: python tools/ncu_footprint.py rep"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(r for r in rows if r and r[0] == "Address")
ci = hdr.index("Instructions Executed"); si = hdr.index("stall_no_inst") if "stall_no_inst" in hdr else None
ex = []
for r in rows:
    if len(r) != len(hdr) or r[0] == "Address": continue
    try: ex.append((int(r[ci]), int(r[si]) if si is not None else 0))
    except ValueError: pass
n = len(ex); tot = sum(e for e, _ in ex); mx = max(e for e, _ in ex)
hot = [e > 0 for e, _ in ex]
print(f"static instructions {n} ({n * 16 / 1024:.1f} KB); executed at least once {sum(hot)} ({sum(hot) * 16 / 1024:.1f} KB)")
for thr in (1e-4, 1e-3, 1e-2, 0.05):
    k = sum(1 for e, _ in ex if e > thr * mx)
    print(f"  executed more than {thr:g} x the hottest instruction: {k} instructions ({k * 16 / 1024:.1f} KB), {100 * sum(e for e, _ in ex if e > thr * mx) / tot:.2f} % of dynamic instructions")
# 128-byte lines (8 instructions) touched by the instructions above 1e-3
lines = {i // 8 for i, (e, _) in enumerate(ex) if e > 1e-3 * mx}
print(f"  128-byte instruction lines holding them: {len(lines)} ({len(lines) * 128 / 1024:.1f} KB)")
# map of the kernel: per 256 instructions, share executed / dynamic weight
print("  map (per 256 static instructions = 4 KB): share of dynamic instructions, instructions executed > 1e-3 x max")
for a in range(0, n, 256):
    seg = ex[a:a + 256]
    print(f"    {a:5d}: {100 * sum(e for e, _ in seg) / tot:5.1f} %   {sum(1 for e, _ in seg if e > 1e-3 * mx):3d}   no_inst samples {sum(s for _, s in seg)}")
