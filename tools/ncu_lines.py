"""Per-CUDA-line summary of an ncu report's source page: ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > X.csv;
python tools/ncu_lines.py X.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hdr = None
lines = {}
for r in rows:
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or r[2] != "-":
        continue
    ie = hdr.index("Instructions Executed"); sa = hdr.index("# Samples")
    wf = hdr.index("L1 Wavefronts Shared") if "L1 Wavefronts Shared" in hdr else None
    wi = hdr.index("L1 Wavefronts Shared Ideal") if wf is not None else None
    try:
        k = (int(r[0]), r[1][:100])
    except ValueError:
        continue
    v = lines.setdefault(k, [0, 0, 0, 0])
    v[0] += int(r[ie] or 0); v[1] += int(r[sa] or 0)
    if wf is not None:
        v[2] += int(r[wf] or 0); v[3] += int(r[wi] or 0)
tot = sum(v[0] for v in lines.values()); ts = sum(v[1] for v in lines.values()); tw = sum(v[2] for v in lines.values())
print("total warp inst %d, samples %d, shared wavefronts %d" % (tot, ts, tw))
for (ln, src), v in sorted(lines.items()):
    if v[0] * 100.0 / tot >= minpct or v[1] * 100.0 / max(ts, 1) >= minpct:
        print("%4d inst %5.2f%% samp %5.2f%% wf %5.2f%% (ideal %5.2f%%) | %s" % (ln, v[0] * 100.0 / tot, v[1] * 100.0 / max(ts, 1), v[2] * 100.0 / max(tw, 1), v[3] * 100.0 / max(tw, 1), src))
