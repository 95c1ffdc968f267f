#!/usr/bin/env python
"""Count the SASS mnemonics that prove the Blackwell paths (tcgen05 = UTCHMMA/UTCBAR/LDTM/UTCATOMSWS, bulk TMA =
UBLKCP, mbarrier = SYNCS, vector reductions = REDG, programmatic dependent launch = ACQBULK (griddepcontrol.wait)
and PREEXIT (griddepcontrol.launch_dependents)) per kernel of the built
library.  Usage: python bench/sass_mnemonics.py > profiles/sass_mnemonics_per_kernel.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributedtf_b200", "lib", "libdtf_b200.so")
KEEP = re.compile(r"^(UTC\w+|LDTM|STTM|UBLKCP|UTMA\w*|SYNCS|REDG|RED|ACQBULK|ARRIVES|CCTL|MEMBAR|ERRBAR|LDG|STG|LD|ST|ATOM\w*)(\.|$)")
INTEREST = re.compile(r"^(UTC|LDTM|STTM|UBLKCP|UTMA|SYNCS|REDG|ACQBULK|PREEXIT|ARRIVES)")
# communication kernels: system-scope release / acquire accesses, fences, 128-bit peer loads, system-scope reductions
COMM_FN = re.compile(r"exploit_copy|allreduce_update|score_|dp_step_begin|head_kernel")
COMM_OPS = re.compile(r"^(LD|ST|LDG|STG|RED|REDG|ATOM|ATOMG|MEMBAR|ERRBAR|FENCE|CCTL)\b.*(SYS|STRONG|128|MEMBAR|ERRBAR|FENCE)")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.Counter()
    fn = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = m.group(1)
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]*)", line)
        if m and fn and INTEREST.match(m.group(1)):
            counts[(fn, m.group(1))] += 1
        elif m and fn and COMM_FN.search(fn) and COMM_OPS.match(m.group(1)):
            counts[(fn, m.group(1))] += 1
    for (fn, op), n in sorted(counts.items()):
        print(n, fn, op)


if __name__ == "__main__":
    sys.exit(main())
