#!/bin/bash
# SASS evidence of the built library (runs without a GPU): per kernel, the Blackwell opcodes that prove the
# tcgen05 / TMEM / TMA paths (B200_PROFILING.md "What proves a Blackwell-native kernel"), register counts and
# spills.  Usage: tools/sass_summary.sh > profiles/r2_sass_opcodes.txt
set -e
LIB=${1:-cslam_b200/_lib/libcslam_nns.so}
echo "# $(basename $LIB): cuobjdump -sass, opcode counts per kernel"
cuobjdump -lelf $LIB | sed 's/^/# /'
echo "# UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = cp.async.bulk.tensor (TMA), UBLKCP = cp.async.bulk,"
echo "# UTCBAR = tcgen05.commit, SYNCS = mbarrier ops; HMMA (legacy mma.sync) must be absent"
cuobjdump -sass $LIB | python3 -c '
import re, sys, collections, subprocess
ops = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "SYNCS", "HMMA", "ELECT", "UCGABAR", "MEMBAR",
       "ATOM", "RED", "LDG", "STG", "LDS", "STS"]
cur, counts = None, collections.OrderedDict()
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); counts[cur] = collections.Counter(); continue
    if cur is None: continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m:
        op = m.group(1)
        counts[cur]["_total"] += 1
        for o in ops:
            if op.startswith(o): counts[cur][o] += 1
def nice(fn):
    name = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
    name = name.replace("(anonymous namespace)::", "").replace("void ", "")
    return re.sub(r"\(.*", "", name)[:64]
print("%-64s %7s  %s" % ("kernel", "instrs", "opcode counts"))
for fn, c in counts.items():
    print("%-64s %7d  %s" % (nice(fn), c["_total"], "  ".join("%s=%d" % (o, c[o]) for o in ops if c[o])))
'
echo
echo "# registers / spills / shared memory per kernel (nvcc -Xptxas -v logs of this build)"
python3 - <<'PY'
import glob, re, subprocess
for f in sorted(glob.glob("cslam_b200/csrc/build/*.ptxas.log")):
    txt = open(f).read()
    for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?Function properties for \S+\n\s*(.*?)\n.*?Used (.*?)\n", txt, re.S):
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "").replace("void ", ""))[:64]
        if name.startswith("cub::"):
            continue
        print("%-64s %s; %s" % (name, m.group(3).strip(), m.group(2).strip()))
PY
