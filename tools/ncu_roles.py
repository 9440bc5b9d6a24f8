#!/usr/bin/env python
"""Per-instruction-block view of one captured launch: consecutive SASS instructions with the same execution count are
one block (a warp role's loop body); prints each block's share of the stall samples and of the issued instructions.
usage: python tools/ncu_roles.py REPORT.ncu-rep [--top N]"""
import csv, io, subprocess, sys


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return 0.0


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index('--top') + 1]) if '--top' in sys.argv else 0
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass', '--kernel-id', ':::1'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if 'Source' in r and any('Sampl' in c for c in r))
    h = rows[hi]
    si, c0, ie = h.index('Source'), h.index('Warp Stall Sampling (All Samples)'), h.index('Instructions Executed')
    body = [r for r in rows[hi + 1:] if len(r) > max(c0, ie, si)]
    body = body[:len(body) // 2] if len(body) > 1 and body[0][si] == body[len(body) // 2][si] else body
    tot = sum(num(r[c0]) for r in body) or 1.0
    tot_i = sum(num(r[ie]) for r in body) or 1.0
    cur, acc, acc_i, start = None, 0.0, 0.0, 0
    for i, r in enumerate(body + [None]):
        e = r[ie] if r else None
        if e != cur:
            if cur is not None and (acc / tot > 0.002 or acc_i / tot_i > 0.002):
                print('%5d-%5d exec/instr=%-10s n=%-4d samples=%6.2f%%  issued=%6.2f%%' % (start, i - 1, cur, i - start, 100 * acc / tot, 100 * acc_i / tot_i))
            cur, acc, acc_i, start = e, 0.0, 0.0, i
        if r:
            acc += num(r[c0]); acc_i += num(r[ie])
    if top:
        idx = sorted(range(len(body)), key=lambda i: -num(body[i][c0]))
        for i in idx[:top]:
            print('%6.2f%% #%d exec=%s %s' % (100 * num(body[i][c0]) / tot, i, body[i][ie], body[i][si][:110]))


if __name__ == '__main__':
    main()
