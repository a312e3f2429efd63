#!/usr/bin/env python
"""Per-kernel SASS evidence for libdli_b200.so: counts of the Blackwell mnemonics that prove tcgen05 / TMEM / TMA use
(UTCHMMA = tcgen05.mma kind::f16, UTCIMMA = kind::i8, LDTM / STTM = tcgen05.ld / .st, UTMALDG = TMA tensor load, UBLKCP = bulk copy, UTCBAR =
tcgen05.commit, HMMA = mma.sync) plus registers and shared memory from the ptxas logs of the build.

    python profiles/sass_summary.py > profiles/r02_sass_summary.txt     (no GPU needed)
"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "distributed_llm_inference_b200", "csrc")
MNEMONICS = ("UTCHMMA", "UTCIMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "HMMA", "LDSM", "RED", "MUFU.EX2", "ACQBULK")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(CSRC, "libdli_b200.so")], capture_output=True, text=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for mn in MNEMONICS:
            if re.search(r"\b" + re.escape(mn) + r"\b", line):
                counts[cur][mn] += 1
    res = {}
    for log in glob.glob(os.path.join(CSRC, "build", "*.log")):
        fn = None
        for line in open(log):
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                fn = m.group(1)
            m = re.search(r"Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", line)
            if m and fn:
                res[fn] = (int(m.group(1)), int(m.group(2) or 0))
            m = re.search(r"(\d+) bytes spill stores", line)
            if m and fn and int(m.group(1)):
                res[fn + "#spill"] = int(m.group(1))
    names = demangle(list(counts))
    print("kernel | regs | static smem B | " + " | ".join(MNEMONICS))
    for fn, c in counts.items():
        short = re.sub(r"\(.*", "", names.get(fn, fn)).replace("dli::", "")
        regs, smem = res.get(fn, ("?", "?"))
        spill = f" SPILL {res[fn + '#spill']}" if fn + "#spill" in res else ""
        print(f"{short} | {regs}{spill} | {smem} | " + " | ".join(str(c.get(mn, 0)) for mn in MNEMONICS))


if __name__ == "__main__":
    main()
