"""Per-kernel counts of the Blackwell-specific SASS mnemonics (proof that tcgen05 / TMEM / TMA / NVLS paths
are what ships).  usage: python scripts/sass_summary.py  -> docs/sass/*.txt"""
import os, re, subprocess, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
OUT = os.path.join(ROOT, "docs", "sass")
os.makedirs(OUT, exist_ok=True)
PAT = re.compile(r"^(UCGABAR_ARV|UCGABAR_WAIT|UTC\w*MMA|UTCBAR|UTCATOM\w*|UTMALDG|UTMASTG|UTMAPF|UBLKCP|LDTM|STTM|SYNCS|LDGSTS|LDGMC|STGMC|REDGMC|"
                 r"REDG|ATOMG|ATOM|LDG\.E\.\d*\.?STRONG\.SYS|STG\.E\.\d*\.?STRONG\.SYS|LD\.E\.\d*\.?STRONG\.SYS|HMMA|FFMA|CCTL)")
for obj in sorted(os.listdir(OBJ)):
    if not obj.endswith(".o"):
        continue
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
    fn, counts = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            fn = fn.replace("(anonymous namespace)::", "").split("(")[0]
            counts[fn] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][\w.]*)", line)
        if m and fn:
            op = m.group(1)
            if PAT.match(op):
                counts[fn][op.split(".")[0] if op.startswith(("FFMA", "SYNCS", "CCTL")) else op] += 1
    with open(os.path.join(OUT, obj.replace(".o", ".txt")), "w") as f:
        f.write(f"# cuobjdump -sass {obj}: counts of tensor-core / TMEM / TMA / barrier / NVLink-related mnemonics per kernel\n")
        for fn, c in counts.items():
            f.write(f"\n{fn}\n")
            for op, n in sorted(c.items()):
                f.write(f"    {n:5d}  {op}\n")
    print(obj, len(counts), "kernels")
