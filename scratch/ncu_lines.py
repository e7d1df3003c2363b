import csv, subprocess, sys, io
def I(x):
    try: return int(x)
    except Exception: return 0
rep, kern = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; cur = None; agg = {}
for r in rows:
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) - 2: continue
    d = dict(zip(hdr, r))
    if r[0] != "":       # a source line summary row
        line = int(r[0]); src = r[1]
        samp = I(d["# Samples"]); inst = I(d["Instructions Executed"])
        ws = I(d.get("L1 Wavefronts Shared")); wsi = I(d.get("L1 Wavefronts Shared Ideal"))
        key = line
        a = agg.setdefault(key, [src, 0, 0, 0, 0])
        a[1] += samp; a[2] += inst; a[3] += ws; a[4] += wsi
tot = sum(a[1] for a in agg.values()); toti = sum(a[2] for a in agg.values())
print("total samples", tot, "warp inst", toti)
for line, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 40]:
    print("%5d %5.1f%% inst %5.1f%% smemwf %8d ideal %8d | %s" % (line, 100 * a[1] / tot, 100 * a[2] / toti, a[3], a[4], a[0].strip()[:110]))
