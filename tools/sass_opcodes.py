"""Opcode histogram of every kernel in libpie_b200.so (cuobjdump -sass): which kernels carry TMA (UTMALDG), tcgen05
(UTCHMMA = bf16, UTCIMMA = int8, UTCBAR, LDTM), mbarrier (SYNCS), legacy tensor-core (HMMA) and cluster (UCGABAR, ST.ASYNC) instructions.
   python tools/sass_opcodes.py > profiles/sass_opcodes_r02.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "proxy_inference_engine_b200", "libpie_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
MARK = ["UTMALDG", "UBLKCP", "UTMAPF", "UTMACCTL", "UTCHMMA", "UTCIMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "SYNCS", "HMMA", "IMMA", "UCGABAR",
        "STAS", "ST.ASYNC", "LDGSTS", "ACQBULK", "PREEXIT", "ELECT", "MUFU", "LOP3", "FFMA", "LDS", "LDG", "STG", "BAR"]
kern, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); kern[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        kern[cur][m.group(1).split(".")[0] if not m.group(1).startswith("ST.ASYNC") else "ST.ASYNC"] += 1
        kern[cur]["__full__" + m.group(1)] += 1
print("kernel | instructions | " + " ".join(MARK))
for k, c in kern.items():
    tot = sum(v for kk, v in c.items() if not kk.startswith("__full__"))
    cells = []
    for mk in MARK:
        n = c.get(mk, 0) if "." not in mk else sum(v for kk, v in c.items() if kk.startswith("__full__" + mk))
        cells.append(f"{mk}={n}" if n else "")
    print(f"{demangle(k)[:110]} | {tot} | " + " ".join(x for x in cells if x))
