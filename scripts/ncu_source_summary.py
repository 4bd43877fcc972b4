#!/usr/bin/env python
"""Source-page summary of one kernel of an .ncu-rep (ncu -i rep --page source --csv): share of the warp-stall samples and of
the executed instructions per 1 KB block of SASS, the barrier waits, and the 25 hottest instructions.  Text for profiles/."""
import collections, csv, io, subprocess, sys

rep = sys.argv[1]
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
print(rows[0][0], rows[0][1])
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
recs = []
for r in rows[2:]:
    try:
        a = int(r[0], 16); s = int(r[ix['# Samples']] or 0); ie = int(r[ix['Instructions Executed']] or 0)
    except Exception:
        continue
    recs.append((a, r[1].strip(), s, ie, {k: int(r[ix[k]] or 0) for k in stalls}))
base = min(x[0] for x in recs)
tot = sum(x[2] for x in recs); toti = sum(x[3] for x in recs)
print(f"warp-stall samples {tot}, instructions executed (source view) {toti}")
agg = collections.Counter()
for x in recs:
    for k, v in x[4].items():
        agg[k] += v
print("stall reasons, % of samples: " + ", ".join(f"{k[6:]} {100 * v / tot:.1f}" for k, v in agg.most_common(10)))
print("\nbarrier waits (mbarrier try_wait) and bulk copies:")
for x in recs:
    if 'TRYWAIT' in x[1] or 'UBLKCP' in x[1]:
        nxt = [y for y in recs if y[0] == x[0] + 16]
        print(f"  {x[0] - base:06x} {x[1][:64]:64s} executed {x[3]:>12d}  samples here+next {100 * (x[2] + (nxt[0][2] if nxt else 0)) / tot:5.2f} %")
print("\nper 1 KB of SASS: % of samples, % of instructions, top stall reasons")
blk = collections.OrderedDict()
for x in recs:
    b = blk.setdefault((x[0] - base) // 0x400, [0, 0, collections.Counter()])
    b[0] += x[2]; b[1] += x[3]
    for k, v in x[4].items():
        b[2][k] += v
for k, b in blk.items():
    if b[0] / tot > 0.004 or b[1] / toti > 0.004:
        print(f"  {k * 0x400:06x}  samples {100 * b[0] / tot:5.1f} %  inst {100 * b[1] / toti:5.1f} %   " +
              " ".join(f"{n[6:]}={100 * v / tot:.1f}" for n, v in b[2].most_common(4)))
print("\n25 hottest instructions:")
for x in sorted(recs, key=lambda t: -t[2])[:25]:
    print(f"  {x[0] - base:06x} {x[1][:70]:70s} {100 * x[2] / tot:5.2f} %  " + " ".join(f"{n[6:]}={v}" for n, v in sorted(x[4].items(), key=lambda t: -t[1])[:2]))
