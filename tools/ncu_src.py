"""Summarise the source page of an .ncu-rep: top SASS lines by stall samples, stall-reason totals, opcode mix.
usage: python tools/ncu_src.py report.ncu-rep [kernel-regex] [top-n]"""
import collections, csv, re, subprocess, sys
rep = sys.argv[1]
kre = sys.argv[2] if len(sys.argv) > 2 else "."
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}; blocks.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None and r: cur["rows"].append(r)
for b in blocks:
    if not re.search(kre, b["name"]): continue
    h = b["hdr"]; ix = {k: i for i, k in enumerate(h)}
    S = ix["# Samples"]; I = ix["Instructions Executed"]; SRC = ix["Source"]
    tot_s = sum(int(r[S]) for r in b["rows"]); tot_i = sum(int(r[I]) for r in b["rows"])
    print("==", b["name"][:100], "| sass lines", len(b["rows"]), "| warp-inst", tot_i, "| samples", tot_s)
    stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    agg = sorted(((sum(int(r[ix[k]]) for r in b["rows"]), k) for k in stalls), reverse=True)
    print("stalls:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot_s, 1)) for v, k in agg[:8]))
    op = collections.Counter()
    for r in b["rows"]:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[SRC])
        op[m.group(2) if m else "?"] += int(r[I])
    print("opcodes:", ", ".join("%s %.1f%%" % (o, 100.0 * c / max(tot_i, 1)) for o, c in op.most_common(14)))
    top = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][S]))[:topn]
    for i in sorted(top):
        r = b["rows"][i]
        why = sorted(((int(r[ix[k]]), k[6:]) for k in stalls), reverse=True)[:2]
        print("%5d %6.2f%%  %-70s %s" % (i, 100.0 * int(r[S]) / max(tot_s, 1), r[SRC].strip()[:70], " ".join("%s:%d" % (k, v) for v, k in why if v)))
