#!/usr/bin/env python
"""Per-kernel SASS opcode summary of libtsb200.so (what proves which kernels are Blackwell-native):
    python profiles/sass_summary.py > profiles/r2_sass_summary.txt
Counts, per kernel, the tcgen05 MMA (UTCHMMA / UTCQMMA ...), TMA (UTMALDG / UTMASTG), TMEM load/store (LDTM / STTM), tcgen05
commit / barrier (UTCBAR), mbarrier (SYNCS), multimem (.MULTIMEM / REDG ... ) and cluster (UCGABAR, MAPA) mnemonics."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'torchsystem_b200', 'libtsb200.so')
PATTERNS = [('tcgen05.mma', r'\bUTC[A-Z]*MMA'), ('tma.load', r'\bUTMALDG'), ('tma.store', r'\bUTMASTG'), ('tmem.ld', r'\bLDTM'),
            ('tmem.st', r'\bSTTM'), ('tcgen05.commit', r'\bUTCBAR'), ('mbarrier', r'\bSYNCS'), ('tmem.alloc', r'\bUTCATOMSWS|\bUTCCP'),
            ('multimem', r'MULTIMEM|\bLDGMC|\bSTGMC'), ('cluster.barrier', r'\bUCGABAR'), ('mapa/dsmem', r'\bMAPA|\.CLUSTER\b'),
            ('cp.async', r'\bLDGSTS'), ('ffma', r'\bFFMA'), ('hmma (legacy)', r'\bHMMA')]


def main():
    out = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            cur = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r'tsb::\(anonymous namespace\)::|tsb::', '', cur)
            cur = re.sub(r'^void ', '', re.sub(r'\(.*', '', cur))
            kernels.setdefault(cur, collections.Counter())
            continue
        if cur is None or '/*' not in line:
            continue
        kernels[cur]['instructions'] += 1
        for name, pat in PATTERNS:
            if re.search(pat, line):
                kernels[cur][name] += 1
    names = [n for n, _ in PATTERNS]
    print(f'# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a), mnemonic counts per kernel')
    print('kernel'.ljust(58) + 'instr'.rjust(7) + ''.join(n.rjust(16) for n in names))
    tot = collections.Counter()
    for k, c in kernels.items():
        print(k[:57].ljust(58) + str(c['instructions']).rjust(7) + ''.join(str(c[n]).rjust(16) for n in names))
        tot.update(c)
    print('TOTAL'.ljust(58) + str(tot['instructions']).rjust(7) + ''.join(str(tot[n]).rjust(16) for n in names))


if __name__ == '__main__':
    sys.exit(main())
