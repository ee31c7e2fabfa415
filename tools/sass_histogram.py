#!/usr/bin/env python
"""Per-kernel histogram of the Blackwell-specific SASS opcodes in libxvector.so (cuobjdump -sass): the evidence that
the hot kernels are tcgen05 / TMEM / TMA code (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA tensor
load/store, UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, FFMA2/FADD2 = packed fp32 pairs).

  python tools/sass_histogram.py > profiles/rNN_sass_opcodes.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tf-kaldi-speaker_b200", "lib", "libxvector.so")
PAT = re.compile(r"\b(UTCHMMA[.\w]*|UTCQMMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UTMAPF[.\w]*|UBLKCP[.\w]*|UTCBAR[.\w]*|"
                 r"UTCATOMSWS[.\w]*|SYNCS[.\w]*|FFMA2|FADD2|FMUL2|HMMA[.\w]*|RED[.\w]*|ATOMG[.\w]*|ELECT[.\w]*|UCGABAR[.\w]*)")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "").replace("xv::", "")
            name = re.sub(r"^void ", "", re.sub(r"\(.*", "", name))
            cur = kernels.setdefault(name, [collections.Counter(), 0])
            continue
        if cur is None or "/*" not in line:
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][\w.]*)", line)
        if not m:
            continue
        cur[1] += 1
        op = m.group(1)
        if PAT.fullmatch(op):
            cur[0][op] += 1
    total = collections.Counter()
    print("# cuobjdump -sass tf-kaldi-speaker_b200/lib/libxvector.so (sm_100a): Blackwell-specific opcodes per kernel")
    print("# kernel | SASS instructions | opcode=count ...")
    for name, (cnt, n) in kernels.items():
        tc = {k: v for k, v in cnt.items() if not k.startswith(("RED", "ATOMG", "ELECT"))}
        total.update(cnt)
        if not tc:
            continue
        print("%s | %d | %s" % (name, n, " ".join("%s=%d" % kv for kv in sorted(cnt.items()))))
    print("# kernels without tcgen05/TMA/mbarrier opcodes (SIMT): %d of %d" % (
        sum(1 for c, _ in kernels.values() if not any(not k.startswith(("RED", "ATOMG", "ELECT")) for k in c)), len(kernels)))
    print("# library totals: " + " ".join("%s=%d" % kv for kv in sorted(total.items())))


if __name__ == "__main__":
    sys.exit(main())
