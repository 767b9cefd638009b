"""Summarise `ncu --page source --csv --print-source cuda,sass` output: per source line (instructions, stall samples)
and per SASS opcode, for the first profiled launch.   usage: python tools/ncu_src.py src.csv [n_sections_per_launch]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
secs, cur, meta = [], None, {}
for r in rows:
    if len(r) == 2:
        meta[r[0]] = r[1]
    if len(r) > 4 and r[0] == "Line No":
        cur = {"hdr": r, "rows": [], "file": meta.get("File Path", "").split("/")[-1]}
        secs.append(cur)
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["rows"].append(r)
files = [s["file"] for s in secs]
n = int(sys.argv[2]) if len(sys.argv) > 2 else (files[1:].index(files[0]) + 1 if files[0] in files[1:] else len(files))


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


line_inst, line_samp, line_txt = collections.Counter(), collections.Counter(), {}
op_inst, op_samp = collections.Counter(), collections.Counter()
for s in secs[:n]:
    h = s["hdr"]
    iE, iS = h.index("Instructions Executed"), h.index("# Samples")
    curline = None
    for r in s["rows"]:
        if r[0]:  # a source line row: following SASS rows belong to it
            curline = (s["file"], r[0])
            line_txt[curline] = r[1].strip()[:110]
        elif r[2] not in ("", "-"):
            op = r[3].split()
            op = [o for o in op if not o.startswith("@")]
            opc = op[0].split(".")[0] if op else "?"
            op_inst[opc] += I(r[iE]); op_samp[opc] += I(r[iS])
            if curline:
                line_inst[curline] += I(r[iE]); line_samp[curline] += I(r[iS])
tot, ts = sum(op_inst.values()), sum(op_samp.values())
print("total warp-instructions", tot, "samples", ts)
print("-- by opcode")
for k, v in op_inst.most_common(25):
    print(f"  {k:12s} inst={v:9d} ({v / tot:5.1%}) samp={op_samp[k]:6d} ({op_samp[k] / max(ts, 1):5.1%})")
print("-- by source line (instructions)")
for k, v in line_inst.most_common(40):
    print(f"  {k[0]:20s}:{k[1]:>4s} inst={v:9d} ({v / tot:5.1%}) samp={line_samp[k]:6d} ({line_samp[k] / max(ts, 1):5.1%})  {line_txt[k]}")
print("-- by source line (stall samples)")
for k, v in line_samp.most_common(25):
    print(f"  {k[0]:20s}:{k[1]:>4s} samp={v:6d} ({v / max(ts, 1):5.1%}) inst={line_inst[k]:9d}  {line_txt[k]}")
