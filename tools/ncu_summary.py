"""Key counters of .ncu-rep files as a markdown table (usage: python tools/ncu_summary.py a.ncu-rep b.ncu-rep ...)."""
import csv, subprocess, sys
WANT = [("gpu__time_duration.sum", "time"), ("sm__cycles_elapsed.avg.per_second", "SM clock"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM read"),
        ("launch__registers_per_thread", "regs/thread"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__cluster_size", "cluster"), ("smsp__inst_executed.sum", "warp instructions")]
cols = []
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u, v = rows[0], rows[1], rows[2]
    ix = {n: i for i, n in enumerate(h)}
    d = {"kernel": v[ix["Kernel Name"]].split("(")[0].replace("void svb::", "")[:60]}
    for key, label in WANT:
        if key in ix:
            d[label] = (v[ix[key]] + " " + u[ix[key]]).strip()
    cols.append(d)
labels = ["kernel"] + [l for _, l in WANT]
print("| metric | " + " | ".join(c["kernel"] for c in cols) + " |")
print("|---|" + "---|" * len(cols))
for l in labels[1:]:
    print("| " + l + " | " + " | ".join(c.get(l, "") for c in cols) + " |")
