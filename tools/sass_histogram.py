"""SASS opcode histogram of libfgdm_b200.so (run in the build container): proves which hardware paths the kernels use.
python tools/sass_histogram.py > profiles/r02_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fgdm_b200", "libfgdm_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "UTCBAR", "SYNCS", "USETMAXREG",
       "MUFU.EX2", "MUFU.TANH", "HMMA", "FENCE.VIEW.ASYNC", "BAR.SYNC", "UTCATOMSWS", "ELECT"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    per, cur = collections.OrderedDict(), None
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m:
            op = m.group(1)
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    per[cur][o] += 1
    names = subprocess.run(["cu++filt"] + list(per), capture_output=True, text=True).stdout.splitlines()
    tot = collections.Counter()
    for c in per.values():
        tot.update(c)
    print("# SASS opcode histogram of fgdm_b200/libfgdm_b200.so (cuobjdump -sass, sm_100a); tools/sass_histogram.py")
    print("# tcgen05.mma -> UTCHMMA, tcgen05.ld/st -> LDTM/STTM, cp.async.bulk.tensor load/store -> UTMALDG/UTMASTG, mbarrier -> SYNCS,")
    print("# setmaxnreg -> USETMAXREG, tcgen05.commit -> UTCBAR; HMMA (mma.sync, the legacy path) must be absent.\n")
    print("total: " + ", ".join(f"{o} {tot[o]}" for o in OPS) + "\n")
    for (fn, c), nm in zip(per.items(), names):
        if not c:
            continue
        nm = re.sub(r"\(.*", "", nm)
        print(f"{nm[:82]:82s} " + " ".join(f"{o}={c[o]}" for o in OPS if c[o]))


if __name__ == "__main__":
    sys.exit(main())
