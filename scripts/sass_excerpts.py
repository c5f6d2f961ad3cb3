"""SASS excerpts of the hot kernels (not mnemonic counts): for every selected kernel of lib/libdcgan_b200.so, the
instructions around each tensor-core / TMA / TMEM / multimem instruction, with context, written to docs/sass/.
usage: python scripts/sass_excerpts.py   (needs cuobjdump; no GPU)"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_tensorflow_for_dcgan_b200", "lib", "libdcgan_b200.so")
OUT = os.path.join(ROOT, "docs", "sass")
KEY = re.compile(r"UTCHMMA|UTCBAR|UTMALDG|UTMASTG|UBLKCP|LDTM|STTM|UTCATOMSWS|SYNCS|LDGMC|STGMC|multimem|REDG|ELECT|UTMAPF|ACQBULK|USETMAXREG|\bRED\.|ATOMG|LD\.E\.STRONG\.SYS|ST\.E\.STRONG\.SYS|CCTL|STRONG\.SYS|MEMBAR")
SELECT = [
    ("igemm2_F_bn128", r"igemm2_kernelILi0ELi128ELi3ELi2ELi1E"),
    ("igemm2_T_bn64", r"igemm2_kernelILi1ELi64ELi4ELi2ELi1E"),
    ("igemm2_W_bn128", r"igemm2_kernelILi2ELi128ELi3ELi2ELi1E"),
    ("igemm2_T_bn64_two_kblocks_per_slot", r"igemm2_kernelILi1ELi64ELi2ELi2ELi2E"),
    ("c3_fprop_tc", r"c3_fprop_tc_kernel"),
    ("c3_wgrad_tc", r"c3_wgrad_tc_kernel"),
    ("c3_tconv_tc", r"c3_tconv_tc_kernel"),
    ("rs_adam_ag", r"rs_adam_ag_kernel"),
    ("async_apply", r"async_apply_kernel"),
    ("allreduce_small", r"allreduce_small_kernel"),
    ("lin_fwd_gather", r"lin_fwd_gather_kernel"),
]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs, cur, name = {}, [], None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if name:
            funcs[name] = cur
        name, cur = m.group(1), []
    elif name is not None:
        cur.append(line)
if name:
    funcs[name] = cur
os.makedirs(OUT, exist_ok=True)
for tag, rx in SELECT:
    match = [n for n in funcs if re.search(rx, n)]
    if not match:
        print("no kernel matches", rx); continue
    n = match[0]
    body = [l for l in funcs[n] if re.search(r"/\*[0-9a-f]{4,6}\*/", l)]
    ins = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in body]
    ins = [l for l in ins if l.strip() and not re.fullmatch(r"\s*/\* 0x[0-9a-f]+ \*/", l)]
    hits = [i for i, l in enumerate(ins) if KEY.search(l)]
    keep = set()
    for h in hits:
        keep.update(range(max(0, h - 4), min(len(ins), h + 5)))
    with open(os.path.join(OUT, f"excerpt_{tag}.txt"), "w") as f:
        f.write(f"# {n}\n# {len(ins)} SASS instructions; shown: every tensor-core / TMA / TMEM / mbarrier / multimem / reduction\n"
                f"# instruction with 4 instructions of context (cuobjdump -sass of lib/libdcgan_b200.so, sm_100a)\n")
        last = -2
        for i in sorted(keep):
            if i != last + 1:
                f.write("        ...\n")
            f.write(ins[i] + "\n")
            last = i
    print(tag, len(ins), "instructions,", len(hits), "key instructions")
