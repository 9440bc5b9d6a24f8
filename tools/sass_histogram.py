#!/usr/bin/env python
"""Per-kernel histogram of the SASS mnemonics that prove a Blackwell-native path (B200_PROFILING.md "What proves a
Blackwell-native kernel"): UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UBLKCP / UTMALDG / UTMASTG (bulk async copies,
TMA), LDGSTS (cp.async), 128-bit global stores (STG.E*.128), HMMA (legacy mma.sync — expected: none).

    python tools/sass_histogram.py [lib.so] > profiles/sass_histogram.txt      (run by __graft_entry__.build())
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "deep-learning-air-hockey-robot_b200", "lib", "libahk.so")
PATTERNS = (("UTC*MMA (tcgen05.mma)", r"\bUTC\w*MMA"), ("LDTM (tcgen05.ld)", r"\bLDTM"), ("STTM (tcgen05.st)", r"\bSTTM"),
            ("UBLKCP (cp.async.bulk)", r"\bUBLKCP"), ("UTMALDG/UTMASTG (TMA tensor)", r"\bUTMA(LDG|STG)"), ("LDGSTS (cp.async)", r"\bLDGSTS"),
            ("STG.*.128", r"\bSTG\.[\w.]*128"), ("LDG.*.128", r"\bLDG\.[\w.]*128"), ("SYNCS (mbarrier)", r"\bSYNCS"),
            ("HMMA (legacy mma.sync)", r"\bHMMA"), ("ACQBULK/griddepcontrol", r"\b(ACQBULK|PREEXIT)"))


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    name, counts, total = None, collections.OrderedDict(), collections.Counter()
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            name = re.sub(r"\(.*", "", name).replace("ahk::policy::", "").replace("ahk::", "").replace("(anonymous namespace)::", "")
            counts[name] = collections.Counter()
            continue
        if name is None or "/*" not in line:
            continue
        counts[name]["instructions"] += 1
        for label, pat in PATTERNS:
            if re.search(pat, line):
                counts[name][label] += 1
                total[label] += 1
    print("# SASS mnemonic histogram of %s (sm_100a), tools/sass_histogram.py" % os.path.relpath(lib, ROOT))
    labels = [p[0] for p in PATTERNS]
    print("%-64s %8s  %s" % ("kernel", "instr", "  ".join("%s" % l for l in labels)))
    for k, c in counts.items():
        if not any(c[l] for l in labels):
            continue
        print("%-64s %8d  %s" % (k[:64], c["instructions"], "  ".join("%*d" % (len(l), c[l]) for l in labels)))
    print("%-64s %8s  %s" % ("TOTAL", "", "  ".join("%*d" % (len(l), total[l]) for l in labels)))


if __name__ == "__main__":
    main()
