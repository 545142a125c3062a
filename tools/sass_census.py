"""SASS census of the built objects (cuobjdump -sass per object): which kernels carry tcgen05 / TMA instructions.
usage: python tools/sass_census.py > profiles/rNN_sass_census.txt   (after __graft_entry__.build())"""
import glob, os, re, subprocess, collections
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "speakerverification_b200", "csrc")
cols = [("UTCHMMA", r"\bUTCHMMA\b(?!\.2CTA)"), ("UTCHMMA.2CTA", r"UTCHMMA\.2CTA"), ("LDTM", r"\bLDTM"),
        ("UTMALDG", r"\bUTMALDG(?!\S*MULTICAST)"), ("UTMALDG.MCAST", r"UTMALDG\S*MULTICAST"), ("UTMASTG", r"\bUTMASTG"),
        ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS")]
print("# SASS census of libsvb (cuobjdump -sass per object; sm_100a)")
print("# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA tensor load / store, UBLKCP = cp.async.bulk (1-D), SYNCS = mbarrier ops")
print("%-16s" % "object" + "".join("%14s" % c for c, _ in cols))
tot = collections.Counter()
rows = []
for o in sorted(glob.glob(os.path.join(root, "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", o], capture_output=True, text=True).stdout
    cnt = [len(re.findall(rx, sass)) for _, rx in cols]
    for (c, _), n in zip(cols, cnt):
        tot[c] += n
    tot["F2FP.SATFINITE.F16"] += len(re.findall(r"F2FP\.SATFINITE\.F16", sass))
    rows.append((sum(cnt), os.path.basename(o), cnt))
for _, name, cnt in sorted(rows, reverse=True):
    print("%-16s" % name + "".join("%14d" % n for n in cnt))
print("\n# whole library")
for k, v in tot.items():
    print("%-22s %d" % (k, v))
