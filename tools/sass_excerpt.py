#!/usr/bin/env python3
"""SASS evidence for profiles/: for the kernels named on the command line (substring of the mangled
name) print the opcode mix and every line that shows the sm_100a features the design relies on:
bulk/TMA copies (UBLKCP), mbarrier waits (SYNCS), shared atomics (ATOMS), votes, programmatic
dependent launch (ACQBULK/PREEXIT-style griddepcontrol lowering), named barriers (BAR).
usage: sass_excerpt.py libsort.so k_onesweepILi256ELi56ELi3ELi1ELi8 k_histogram_lanesILi4 ..."""
import collections
import re
import subprocess
import sys

KEEP = re.compile(r"\b(UBLKCP|UTMA\w*|SYNCS|ATOMS|ATOMG|REDG?|VOTE|MATCH|ACQBULK|PREEXIT|BAR|UCGABAR\w*|FENCE|MEMBAR|CCTL|ERRBAR)\b")


def main():
    lib, pats = sys.argv[1], sys.argv[2:]
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    fn, body = None, collections.OrderedDict()
    for line in out.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            body[fn] = []
            continue
        if fn and re.match(r"\s+/\*[0-9a-f]{4,5}\*/", line):
            body[fn].append(line.rstrip())
    for f, lines in body.items():
        if not any(p in f for p in pats):
            continue
        demangled = subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip()
        mix = collections.Counter()
        for l in lines:
            m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
            if m:
                mix[m.group(2)] += 1
        print("=" * 100)
        print(demangled[:300])
        print("SASS instructions: %d ; mix: %s" % (sum(mix.values()), ", ".join("%s %d" % kv for kv in mix.most_common(18))))
        seen = collections.Counter()
        for l in lines:
            m = KEEP.search(l)
            if m:
                seen[m.group(1)] += 1
                if seen[m.group(1)] <= 4:                 # first few of each kind, with their operands
                    print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l))
        print("feature lines:", dict(seen))


if __name__ == "__main__":
    main()
