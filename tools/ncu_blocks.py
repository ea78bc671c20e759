"""Hot basic blocks per kernel from an `ncu --page source --csv --print-source sass` dump:
python tools/ncu_blocks.py gpurun_out/ncu_x_source_sass.csv [top_n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 8
kern, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        kern.append(cur)
        continue
    if r and r[0] == "Address":
        cur["hdr"] = r
        continue
    if cur is not None and r:
        cur["rows"].append(r)
for k in kern:
    h = k["hdr"]
    isrc, ie, ismp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    ite = h.index("Thread Instructions Executed")
    tot = sum(float(r[ie] or 0) for r in k["rows"])
    tots = sum(float(r[ismp] or 0) for r in k["rows"]) or 1.0
    tt = sum(float(r[ite] or 0) for r in k["rows"])
    print(f"== {k['name'][:70]}  warp-inst {tot:.3e}  avg active lanes {tt / max(tot, 1):.1f}  samples {tots:.0f}")
    byop = collections.Counter()
    bys = collections.Counter()
    for r in k["rows"]:
        op = r[isrc].split()[0] if not r[isrc].startswith("@") else r[isrc].split()[1]
        byop[op.split(".")[0]] += float(r[ie] or 0)
        bys[op.split(".")[0]] += float(r[ismp] or 0)
    print("   inst mix:", {o: round(c / tot * 100, 1) for o, c in byop.most_common(10)})
    print("   stall samples by op:", {o: round(c / tots * 100, 1) for o, c in bys.most_common(8)})
    for name in ["stall_long_sb", "stall_math", "stall_wait", "stall_short_sb", "stall_not_selected", "stall_mio", "stall_lg",
                 "stall_branch_resolving", "stall_dispatch", "stall_selected", "stall_barrier", "stall_membar"]:
        if name in h:
            j = h.index(name)
            t = sum(float(r[j] or 0) for r in k["rows"])
            if t / tots > 0.03:
                print(f"   {name}: {t / tots * 100:.1f}%", end="")
    print()
    blocks = []
    for i, r in enumerate(k["rows"]):
        e, s_ = float(r[ie] or 0), float(r[ismp] or 0)
        op = r[isrc].split()[0] if not r[isrc].startswith("@") else r[isrc].split()[1]
        if blocks and abs(blocks[-1]["e"] - e) < 1e-9:
            blocks[-1]["n"] += 1; blocks[-1]["s"] += s_; blocks[-1]["ops"].append(op)
        else:
            blocks.append({"start": i, "e": e, "n": 1, "s": s_, "ops": [op]})
    blocks.sort(key=lambda b: -b["s"])
    for b in blocks[:top]:
        c = collections.Counter(o.split(".")[0] for o in b["ops"])
        print(f"   sass[{b['start']:5d}+{b['n']:3d}] exec {b['e']:.2e} inst-share {b['e'] * b['n'] / tot * 100:5.1f}% "
              f"samples {b['s'] / tots * 100:5.1f}%  {dict(c.most_common(7))}")
