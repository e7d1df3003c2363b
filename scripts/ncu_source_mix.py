"""Opcode mix, stall samples and shared-memory wavefronts per opcode (and the hottest lines) from
`ncu -i REP --page source --csv --kernel-name regex:NAME --launch-count 1 > src.csv`.  Usage: ncu_source_mix.py src.csv [top_lines]"""
import csv, collections, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; idx = {h: i for i, h in enumerate(hdr)}
ops = collections.defaultdict(lambda: [0, 0, 0]); tot_s = tot_i = 0; data = []
def num(x):
    try: return int(float(x))
    except Exception: return 0
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[0] == "Address": continue
    src = r[idx["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2) if m else src
    base = op.split(".")[0]
    key = ".".join(op.split(".")[:3]) if base in ("MUFU", "F2F", "I2F", "F2I", "LDS", "STS", "LDG", "STG", "LD", "ST") else base
    samp, ex, wf = num(r[idx["# Samples"]]), num(r[idx["Instructions Executed"]]), num(r[idx["L1 Wavefronts Shared"]])
    ops[key][0] += ex; ops[key][1] += samp; ops[key][2] += wf; tot_s += samp; tot_i += ex
    data.append((samp, ex, wf, src))
print(rows[0][1] if len(rows[0]) > 1 else "", "| warp instructions", tot_i, "| stall samples", tot_s)
for k, (e, sm, wf) in sorted(ops.items(), key=lambda x: -x[1][1])[:34]:
    print("%-18s exec %9d (%4.1f%%)  samples %5d (%4.1f%%)  smem wavefronts %8d" % (k, e, 100 * e / max(tot_i, 1), sm, 100 * sm / max(tot_s, 1), wf))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for samp, ex, wf, src in sorted(data, reverse=True)[:top]:
    print("%5d samples  exec %7d  wf %7d  %s" % (samp, ex, wf, src[:110]))
