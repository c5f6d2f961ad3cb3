"""Per-kernel summary table from an `ncu --page raw --csv` export of one training step.
usage: ncu_summary.py raw.csv > profiles/ncu_step_summary.md"""
import csv, re, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, body = rows[0], [r for r in rows[2:] if len(r) == len(rows[0])]
col = {h: i for i, h in enumerate(hdr)}
def f(r, name, scale=1.0):
    try:
        return float(r[col[name]].replace(",", "")) * scale
    except Exception:
        return float("nan")
print("| # | kernel | grid | cluster | regs | smem/CTA KB | us | tensor pipe % | L2 thr % | DRAM thr % | L1 thr % | warps active % | DRAM MB r/w |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
tot = 0.0
for i, r in enumerate(body):
    name = re.sub(r"^void |dcgan::|\(.*", "", r[col["Kernel Name"]])
    name = re.sub(r"at::native::.*", "at::fill (grad memset)", name)[:44]
    us = f(r, "gpu__time_duration.sum")
    tot += us
    print(f"| {i} | `{name}` | {r[col['Grid Size']].strip('()').replace(' ', '')} | {int(f(r, 'launch__cluster_size'))} | "
          f"{int(f(r, 'launch__registers_per_thread'))} | {f(r, 'launch__shared_mem_per_block_dynamic'):.0f} | {us:.1f} | "
          f"{f(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{f(r, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{f(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{f(r, 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{f(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{f(r, 'dram__bytes_read.sum'):.1f}/{f(r, 'dram__bytes_write.sum'):.1f} |")
print(f"\nsum of serialized kernel durations (cold caches between kernels under ncu): {tot:.0f} us over {len(body)} launches")
