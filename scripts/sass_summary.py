#!/usr/bin/env python
"""Per-kernel SASS instruction summary of libkeraslm_b200.so (what proves the Blackwell-native path, B200_PROFILING.md):
UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier,
HMMA = legacy mma.sync (must be absent), LDL / STL = local memory (spills, dynamically indexed arrays).
usage: python scripts/sass_summary.py > profiles/<round>_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, 'ocrd_keraslm_b200', 'libkeraslm_b200.so')
out = subprocess.run(['cuobjdump', '-sass', lib], stdout=subprocess.PIPE, universal_newlines=True, check=True).stdout
watch = ['UTCHMMA', 'UTCQMMA', 'LDTM', 'STTM', 'UBLKCP', 'UTMALDG', 'UTCBAR', 'SYNCS', 'HMMA', 'LDL', 'STL', 'MUFU', 'FFMA', 'SHFL',
         'MEMBAR', 'CCTL', 'ATOM', 'RED', 'BAR']
counts, total, name = collections.OrderedDict(), collections.Counter(), None
for line in out.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        name = subprocess.run(['c++filt', m.group(1)], stdout=subprocess.PIPE, universal_newlines=True).stdout.strip()
        name = re.sub(r'\(.*', '', name)
        counts[name] = collections.Counter()
        continue
    m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
    if m and name:
        op = m.group(1)
        total[name] += 1
        for w in watch:
            if op.startswith(w):
                counts[name][w] += 1
print('SASS summary of %s (%d bytes)' % (os.path.relpath(lib, ROOT), os.path.getsize(lib)))
print('%-46s %7s ' % ('kernel', 'instr') + ' '.join('%7s' % w for w in watch))
for k, c in counts.items():
    print('%-46s %7d ' % (k[:46], total[k]) + ' '.join('%7d' % c[w] for w in watch))
