"""Per-kernel counts of the SASS mnemonics that show how each kernel of libb2c.so moves data and cooperates:
UBLKCP (TMA bulk copy, cp.async.bulk), SYNCS (mbarrier), LDGSTS (cp.async), 128-bit global loads/stores, shared-memory
loads/stores, warp votes / shuffles / matches, and the tensor-core mnemonics (none expected: nothing here is a
contraction).   python tools/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "brpc-github-actions_b200", "libb2c.so")
PAT = collections.OrderedDict([
    ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("LDGSTS", r"\bLDGSTS"), ("LDG.128", r"\bLDG\.E\S*\.128"),
    ("STG.128", r"\bSTG\.E\S*\.128"), ("LDS", r"\bLDS"), ("STS", r"\bSTS"), ("VOTE", r"\bVOTE"), ("SHFL", r"\bSHFL"),
    ("MATCH", r"\bMATCH"), ("REDUX", r"\bREDUX"), ("tensor (HMMA/UTC*MMA)", r"\b(HMMA|UTC\w*MMA|LDTM|STTM)"),
])
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
cur, counts, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur] = collections.Counter()
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        total[cur] += 1
        for k, p in PAT.items():
            if re.search(p, line):
                counts[cur][k] += 1
print("# SASS mnemonic counts per kernel of libb2c.so (cuobjdump -sass, sm_100a); tools/sass_summary.py")
print("%-58s %6s " % ("kernel", "instr") + " ".join("%8s" % k[:8] for k in PAT))
for k, c in counts.items():
    print("%-58s %6d " % (k[-58:], total[k]) + " ".join("%8d" % c[p] for p in PAT))
