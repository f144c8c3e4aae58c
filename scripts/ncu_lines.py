"""Per source line totals from an ncu report with --import-source on: python scripts/ncu_lines.py rep [top]
Prints warp instructions executed, average active threads and stall samples per CUDA source line."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None; hdr = None; rows = []
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and r[2] == "-" and r[0].isdigit():  # a source line row (sass rows have an address)
        d = dict(zip(hdr[4:], r[4:]))
        try:
            inst = int(d["Instructions Executed"]); thr = int(d["Thread Instructions Executed"]); smp = int(d["# Samples"])
        except Exception:
            continue
        rows.append((inst, thr, smp, cur, int(r[0]), r[1].strip()))
tot = sum(x[0] for x in rows); tots = sum(x[2] for x in rows)
print(f"total warp instructions {tot:,}  samples {tots:,}")
byfile = {}
for inst, thr, smp, f, ln, src in rows:
    a = byfile.setdefault(f, [0, 0, 0]); a[0] += inst; a[1] += thr; a[2] += smp
for f, a in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:24s} inst {100*a[0]/tot:5.1f}%  samples {100*a[2]/max(1,tots):5.1f}%  avg threads {a[1]/max(1,a[0]):5.1f}")
for inst, thr, smp, f, ln, src in sorted(rows, reverse=True)[:top]:
    print(f"{100*inst/tot:5.2f}% smp {100*smp/max(1,tots):5.2f}% thr {thr/max(1,inst):5.1f}  {f}:{ln}  {src[:110]}")
# optional: totals for line ranges of one file:  ... rep top file:lo-hi file:lo-hi
for spec in sys.argv[3:]:
    f, rng = spec.split(":"); lo, hi = (int(x) for x in rng.split("-"))
    sel = [x for x in rows if x[3] == f and lo <= x[4] <= hi]
    i = sum(x[0] for x in sel); t = sum(x[1] for x in sel); s_ = sum(x[2] for x in sel)
    print(f"{spec}: inst {100*i/tot:5.1f}%  samples {100*s_/max(1,tots):5.1f}%  avg threads {t/max(1,i):5.1f}")
