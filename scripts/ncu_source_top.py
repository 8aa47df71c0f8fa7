#!/usr/bin/env python3
"""Top SASS instructions by warp-stall samples from `ncu -i X.ncu-rep --page source --csv` (one kernel launch)."""
import csv
import sys


def main(path, n=45):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    data = []
    for r in rows[hi + 1:]:
        if not r or not r[0].startswith('0x'):
            break  # next launch
        if len(r) == len(hdr):
            data.append(r)
    ix = {h: i for i, h in enumerate(hdr)}
    tot = sum(int(r[ix['# Samples']]) for r in data)
    totinst = sum(int(r[ix['Instructions Executed']]) for r in data)
    thr = sum(int(r[ix['Thread Instructions Executed']]) for r in data)
    print('samples', tot, 'warp instructions', totinst, 'avg active threads %.2f' % (thr / totinst), 'SASS lines', len(data))
    stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    agg = {s: sum(int(r[ix[s]]) for r in data) for s in stalls}
    print({k[6:]: round(v / tot, 3) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:9]})
    top = sorted(range(len(data)), key=lambda i: -int(data[i][ix['# Samples']]))[:n]
    for i in sorted(top):
        r = data[i]
        st = sorted(((int(r[ix[s]]), s) for s in stalls), reverse=True)[:2]
        print('%4d %-64s smp %5.2f%% exec %10s thr %5s  %s' % (i, r[ix['Source']].strip()[:64], 100 * int(r[ix['# Samples']]) / tot,
              r[ix['Instructions Executed']], r[ix['Avg. Threads Executed']][:5], ' '.join('%s:%d' % (s[6:], c) for c, s in st)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)
