#!/usr/bin/env python
"""Writes profiles/r02_sass_grep.txt: per kernel of the library, the number of SASS lines with the
mnemonics that show which hardware features it uses.

    python benchmarks/scripts/sass_grep.py > profiles/r02_sass_grep.txt
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, 'scalable_agent_b200', 'lib', 'libvtrace_b200.so')
COLUMNS = [('UBLKCP', r'\bUBLKCP'), ('SYNCS', r'\bSYNCS'), ('UCGABAR', r'\bUCGABAR'), ('ATOM', r'\b(ATOMG|REDG)'),
           ('STL/LDL', r'\b(STL|LDL)\b'), ('MUFU.EX2', r'MUFU\.EX2'), ('SHFL', r'\bSHFL'),
           ('UTCHMMA', r'\bUTC\w*MMA'), ('LDTM', r'\bLDTM'), ('STTM', r'\bSTTM'), ('UTMALDG', r'\bUTMALDG'),
           ('UTCBAR', r'\bUTCBAR'), ('ACQBULK', r'\bACQBULK'), ('PREEXIT', r'\bPREEXIT')]

HEADER = '''# cuobjdump -sass scalable_agent_b200/lib/libvtrace_b200.so, final round-2 build.  Per kernel, the number of SASS lines with:
#   UBLKCP    cp.async.bulk (1-D TMA bulk copy, both directions)      SYNCS     mbarrier operations
#   UCGABAR   barrier.cluster arrive / wait (thread-block clusters; K5 also has S2R SR_CgaCtaId and CGAERRBAR;
#             its st.shared::cluster stores compile to generic stores through the cluster window)
#   ATOM      ATOMG / REDG: the tickets (no value is ever accumulated atomically)
#   STL/LDL   local memory: 0 everywhere = no spills             MUFU.EX2  ex2.approx     SHFL  warp shuffles
#   UTCHMMA   tcgen05.mma      LDTM / STTM  tcgen05.ld / tcgen05.st      UTMALDG  cp.async.bulk.tensor (tensor-map TMA)
#   UTCBAR    tcgen05.commit
#   ACQBULK   griddepcontrol.wait      PREEXIT  griddepcontrol.launch_dependents   (programmatic dependent launch)
# The V-trace / loss kernels have no UTC*MMA / HMMA: nothing on that path is a contraction.  The heads kernel
# (vtr::head::policy_head_kernel) is the one contraction next to it and the one tensor-core kernel.'''


def main():
  sass = subprocess.run(['cuobjdump', '-sass', LIB], check=True, capture_output=True, text=True).stdout
  counts = collections.OrderedDict()
  name = None
  for line in sass.splitlines():
    m = re.match(r'\s*Function : (\S+)', line)
    if m:
      name = m.group(1)
      counts[name] = collections.Counter()
      continue
    if name is None or not re.match(r'\s+/\*[0-9a-f]{4,}\*/', line):
      continue
    counts[name]['instrs'] += 1
    for col, pat in COLUMNS:
      if re.search(pat, line):
        counts[name][col] += 1
  print(HEADER)
  cols = [c for c, _ in COLUMNS] + ['instrs']
  for name in sorted(counts):
    print('%-76s' % name + ' '.join('%*d' % (max(5, len(c)), counts[name][c]) for c in cols))
  print('%-76s' % 'kernel' + ' '.join('%*s' % (max(5, len(c)), c) for c in cols))


if __name__ == '__main__':
  main()
