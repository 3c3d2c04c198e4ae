"""Per-kernel counts of the SASS mnemonics that show what a kernel runs on (B200_PROFILING.md, "What proves a
Blackwell-native kernel"): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA tensor load,
UBLKCP/UBLKRED = cp.async.bulk / cp.reduce.async.bulk, SYNCS = mbarrier, FFMA/IMAD/HMMA for contrast.
  python scripts/sass_summary.py caltech-bifrost-dsp_b200/xengine/libxeng.so > profiles/r02_sass_summary.txt"""
import collections
import re
import subprocess
import sys

so = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
names = subprocess.run(["cu++filt"], input=txt, capture_output=True, text=True).stdout if False else txt
WATCH = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKRED", "SYNCS",
         "HMMA", "IMMA", "FFMA", "IMAD", "LDG", "STG", "LDS", "STS", "RED", "ATOM", "REDUX", "ELECT"]
cur, counts, total = None, collections.OrderedDict(), {}
for ln in txt.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        total[cur] = 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if m and cur:
        total[cur] += 1
        op = m.group(1)
        for w in WATCH:
            if op.startswith(w):
                counts[cur][w] += 1
                break
dem = subprocess.run(["c++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
print("source: cuobjdump -sass %s (sm_100a)" % so)
print("kernel | instructions | " + " ".join(WATCH))
for k, d in zip(counts, dem):
    short = re.sub(r"\(anonymous namespace\)::|xeng::", "", d)[:70]
    print("%s | %d | %s" % (short, total[k], " ".join("%s=%d" % (w, counts[k][w]) for w in WATCH if counts[k][w])))
