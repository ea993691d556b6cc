#!/usr/bin/env python
"""Per-kernel counts of the SASS instructions that prove the tcgen05 / TMEM / TMA path, from the
built library:  python tools/sass_summary.py > profiles/r02_sass_summary.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'rnb_b200', 'csrc', 'librnb_b200.so')
MNEMONICS = ['UTCHMMA', 'UTCBAR', 'UTMALDG', 'UTMASTG', 'UTMAPF', 'LDTM', 'STTM', 'UTCATOM',
             'SYNCS', 'ACQBULK', 'UCGABAR', 'HMMA']


def main():
    out = subprocess.run(['cuobjdump', '-sass', LIB], stdout=subprocess.PIPE, text=True,
                         check=True).stdout
    arch = sorted(set(re.findall(r'arch = (sm_\w+)', out)))
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r'\s*Function : (\S+)', line)
        if m:
            cur = subprocess.run(['c++filt', m.group(1)], stdout=subprocess.PIPE,
                                 text=True).stdout.strip()
            cur = cur.replace('(anonymous namespace)::', '').replace('void ', '').replace('rnb::', '')
            cur = re.sub(r'\(.*', '', cur)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r'/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', line)
        if m:
            op = m.group(1).split('.')[0]
            counts[cur][op] += 1
            counts[cur]['_total'] += 1
    print('# SASS summary of rnb_b200/csrc/librnb_b200.so (%s)\n' % ', '.join(arch))
    print('`cuobjdump -sass`, instruction counts per kernel.  UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld,')
    print('UTMALDG = TMA tensor load, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops.\n')
    print('| kernel | instructions | ' + ' | '.join(MNEMONICS) + ' |')
    print('|---|---|' + '---|' * len(MNEMONICS))
    tot = collections.Counter()
    for k, c in counts.items():
        print('| `%s` | %d | %s |' % (k, c['_total'], ' | '.join(str(c[m]) for m in MNEMONICS)))
        tot.update(c)
    print('| **all kernels** | %d | %s |' % (tot['_total'], ' | '.join(str(tot[m]) for m in MNEMONICS)))


if __name__ == '__main__':
    main()
