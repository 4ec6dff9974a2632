"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: one device step, by kernel and grid."""
import collections
import csv
import sys


def main(path, out=None, note=''):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rows = [(r['Kernel Name'], r['Grid Size'], float(r['Metric Value'])) for r in csv.DictReader(lines)]
    idx = [i for i, (k, g, t) in enumerate(rows) if 'k_blur_axis<float, 0>' in k or 'k_blur_rows' in k or 'k_blur_tile' in k]
    if not idx:
        s, e = 0, len(rows)
    else:
        s = idx[-2] if len(idx) >= 2 and idx[-1] - idx[-2] > 200 else idx[-1]
        e = idx[-1] if s != idx[-1] else len(rows)
    # a device step of bench.py ends with the epilogue; what follows (the smoother-only roofline loop) is not the step
    ends = [i for i, (k, g, t) in enumerate(rows) if 'k_epilogue' in k and s <= i < e]
    if ends:
        e = ends[0] + 1
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, g, t in rows[s:e]:
        k = k.split('(ofb::')[0].split('(long')[0].split('(int')[0].split('(const')[0]
        agg[(k[:90], g)][0] += 1
        agg[(k[:90], g)][1] += t
    tot = sum(v[1] for v in agg.values())
    text = ['%s' % note, 'one device step: %d launches, sum of kernel time %.2f ms' % (e - s, tot / 1e6)]
    for (k, g), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
        text.append('%9.1f us %5.1f%% n=%4d avg=%7.1fus grid=%-18s %s' % (v[1] / 1e3, 100 * v[1] / tot, v[0], v[1] / v[0] / 1e3, g, k))
    text = '\n'.join(text)
    print(text)
    if out:
        open(out, 'w').write(text + '\n')


if __name__ == '__main__':
    main(*sys.argv[1:])
