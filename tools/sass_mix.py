#!/usr/bin/env python3
"""Opcode histogram per kernel from `cuobjdump -sass` (developer tool)."""
import collections, re, subprocess, sys
lib = sys.argv[1]; pat = sys.argv[2] if len(sys.argv) > 2 else ''
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
fn = None; mix = collections.defaultdict(collections.Counter)
for line in out.split('\n'):
    m = re.search(r'Function : (\S+)', line)
    if m: fn = m.group(1); continue
    m = re.match(r'\s+/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
    if m and fn: mix[fn][m.group(2)] += 1
for f, c in mix.items():
    if pat in f:
        print(f, 'total', sum(c.values())); print('   ', c.most_common(24))
