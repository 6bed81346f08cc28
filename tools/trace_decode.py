#!/usr/bin/env python
"""Decode gpurun_out/trace_<name>.txt written by AI4E_B200_PIPE_TRACE (a -DAI4E_PIPE_TRACE build)."""
import sys
for name in sys.argv[1:]:
    rows = [l.split() for l in open(f'gpurun_out/trace_{name}.txt')]
    E = [[int(v) for v in r[2:]] for r in rows if r[0] == 'E']
    T = [[int(v) for v in r[1:]] for r in rows if r[0] == 'T']
    K = [[int(v) for v in r[1:]] for r in rows if r[0] not in ('E', 'T')]
    t0 = min(v for r in K + E for v in r if v > 0)
    print('==', name)
    if T:
        t = T[0]
        print(' phases (cycles from kernel entry): setup %d, prev-kernel wait %d, MMAs issued %d, epilogue done %d, all done %d' % tuple(v - t[0] for v in t[1:]))
        t0 = min(t0, t[0])
    print(' K blocks (slot free, A issued, data seen, MMAs issued):')
    for i, r in enumerate(K[:12]):
        if r[0]: print('  ', i, [v - t0 if v else 0 for v in r])
    prev = None
    for i, r in enumerate(E[:16]):
        if not r[0]: break
        d = [r[j] - r[0] for j in range(len(r))]
        print('  E', i, 'start', r[0] - t0, 'd', d[1:], 'period', (r[0] - prev) if prev else '-')
        prev = r[0]
