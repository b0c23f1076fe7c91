#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of the render path: totals per kernel and, with -r, per round.
usage: python scripts/launch_rounds.py <launches.csv> [-r]"""
import csv, re, sys


def main():
    with open(sys.argv[1]) as f:
        lines = [l for l in f if l.startswith('"')]
    seq = []
    for x in csv.DictReader(lines):
        name = re.sub(r'<.*', '', re.sub(r'\(.*', '', x['Kernel Name']).replace('ngp::', '').replace('void ', ''))
        seq.append((name, float(x['Metric Value'].replace(',', '')) / 1000, x['Grid Size']))
    tot = {}
    for n, us, g in seq:
        tot.setdefault(n, [0.0, 0])
        tot[n][0] += us
        tot[n][1] += 1
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0]):
        print(f"{k:40s} {v[0]:10.1f} us  x{v[1]}")
    if "-r" in sys.argv:
        rnd, cur = [], None
        for n, us, g in seq:
            if n.startswith('compact_rays'):
                if cur: rnd.append(cur)
                cur = {'compact': (us, g)}
            elif cur is not None and n[:4] in ('gene', 'hash', 'mlp_', 'comp'):
                cur[n[:4]] = (us, g)
        if cur: rnd.append(cur)
        for i, r in enumerate(rnd):
            print(i, ' '.join(f"{k}={v[0]:.1f}{v[1]}" for k, v in r.items()))


if __name__ == "__main__":
    main()
