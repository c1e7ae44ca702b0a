#!/usr/bin/env python
"""SASS census of libkubedl_b200.so: per kernel, how often each instruction that proves a Blackwell-native
path (or a cross-GPU one) occurs.  `python scripts/sass_census.py > profiles/sass_instruction_census.txt`."""
import collections
import re
import subprocess
import sys

sys.path.insert(0, ".")
from kubedl_b200.ops import build as B

PAT = re.compile(r"\b(UTC[A-Z]*MMA(?:\.2CTA)?|LDTM(?:\.x\d+)?|STTM|UTMALDG(?:\.\dD)?(?:\.IM2COL)?|UTMASTG(?:\.\dD)?|UTMAPF(?:\.L2)?(?:\.\dD)?|"
                 r"UTMAREDG(?:\.\dD)?|UBLKCP(?:\.[A-Z]+)*|UBLKRED(?:\.[A-Z]+)*|UTCBAR(?:\.2CTA)?(?:\.MULTICAST)?|"
                 r"MULTIMEM(?:\.[A-Z0-9_]+)*|LDGMC(?:\.[A-Z0-9_]+)*|STGMC(?:\.[A-Z0-9_]+)*|REDGMC(?:\.[A-Z0-9_]+)*|REDG?(?:\.[A-Z0-9_]+)*|ATOMG(?:\.[A-Z0-9_]+)*|ATOMS(?:\.[A-Z0-9_]+)*|"
                 r"SYNCS(?:\.[A-Z0-9_]+)*|HMMA(?:\.[A-Z0-9_]+)*)\b")


def main():
    lib = B.LIB_PATH
    out = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
    demangle = {}
    names = sorted(set(re.findall(r"Function : (\S+)", out)))
    dm = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines() if names else []
    for n, d in zip(names, dm):
        demangle[n] = re.sub(r"\((?:[^()]|\([^()]*\))*\)$", "", d.strip())
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = demangle.get(m.group(1), m.group(1))
            counts.setdefault(cur, collections.Counter())
            continue
        if cur is None or "/*" not in line:
            continue
        body = line.split("*/", 1)[-1]
        m = PAT.search(body)
        if m:
            op = m.group(1)
            if op.startswith("SYNCS") and "TRANS64" not in op and "EXCH" not in op:
                continue
            counts[cur][op] += 1
    print("# SASS census of libkubedl_b200.so (cuobjdump -sass): count, kernel, mnemonic.")
    print("# UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA tensor load/store (.IM2COL = im2col mode),")
    print("# UTMAPF = TMA prefetch to L2, UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS.*TRANS64 = mbarrier with transaction count,")
    print("# LDGMC / STGMC = NVLS multimem.ld_reduce / multimem.st, RED/REDG/ATOMG = global reductions (peer memory in the exchange kernels), ATOMS = shared atomics.")
    print("# No HMMA (mma.sync / wmma) anywhere.")
    for k, c in counts.items():
        for op, n in sorted(c.items()):
            print(f"{n} {k.replace('(int)', '')} {op}")


if __name__ == "__main__":
    main()
