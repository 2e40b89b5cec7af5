"""List, per kernel of the shipped library, the SASS mnemonics that prove the Blackwell path: UTCHMMA (tcgen05.mma), UTMALDG (TMA load),
LDTM / STTM (tcgen05.ld / st), UTCBAR, and the legacy HMMA (mma.sync).  Runs on the build machine (cuobjdump, no GPU):

    python probes/sass_tensor_ops.py > profiles/sass_tensor_ops_r03.txt
"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "objcavit_b200", "libobjcavit_b200.so")
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
pat = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCOMMA|UTMALDG|UTMASTG|LDTM|STTM|UTCBAR|UTCCP|HMMA|SYNCS)\b")
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "").split("(")[0]
        counts[cur] = collections.Counter()
        continue
    if cur:
        for op in pat.findall(line):
            counts[cur][op] += 1
print(f"cuobjdump -sass {os.path.relpath(LIB, ROOT)}: tensor / TMA / TMEM mnemonics per kernel (kernels with none of them omitted)")
for k, c in counts.items():
    if any(op != "SYNCS" for op in c):
        print(f"{k:<70s} " + "  ".join(f"{op} {n}" for op, n in sorted(c.items())))
