"""Summarise scripts/build/ext_trace output (development aid): python scripts/trace_report.py gpurun_out/trace.log"""
import re
import sys

roles, cur = {}, None
for line in open(sys.argv[1]):
    if line.startswith('=='):
        cur = line[3:].split(':')[0].strip()
        roles[cur] = {}
    elif cur and re.match(r'\s*\d+:', line):
        j, rest = line.split(':')
        roles[cur][int(j)] = [int(x) for x in rest.split()]
names = list(roles)
sm = [n for n in names if n.startswith('softmax')]
js = sorted(roles[sm[0]])
mid = js[6:-4]
d = [roles[sm[0]][j + 1][0] - roles[sm[0]][j][0] for j in mid]
print('steps', len(js), 'steady step clk mean %.0f min %d max %d' % (sum(d) / len(d), min(d), max(d)))
for n in sm:
    R = roles[n]
    ph = [[R[j][i + 1] - R[j][i] for i in range(4)] for j in mid]
    print('%-16s ld %4d  max %4d  exp0 %4d  exp1 %4d  total %4d   idle->next s_full %4d' % (
        n, *[round(sum(p[i] for p in ph) / len(ph)) for i in range(4)],
        round(sum(R[j][4] - R[j][0] for j in mid) / len(mid)), round(sum(R[j + 1][0] - R[j][4] for j in mid) / len(mid))))
M = roles[[n for n in names if n.startswith('mma')][0]]
dm = [[M[j][i + 1] - M[j][i] for i in range(6)] + [M[j + 1][0] - M[j][6]] for j in mid]
print('mma deltas Vrdy->pA0->pA1->S_A->pB0->pB1->S_B->next', [round(sum(x[i] for x in dm) / len(dm)) for i in range(7)])
j = mid[3]
print('one step j=%d (relative to softmax w0 s_full)' % j)
b = roles[sm[0]][j][0]
for n in names:
    if j in roles[n]:
        print('  %-40s' % n[:40], [v - b if v >= 0 else None for v in roles[n][j]])
