"""Per-kernel summary (markdown) of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
python tools/ncu_summary.py profiles/<list>.csv "title" > profiles/<list>_summary.md"""
import collections
import csv
import re
import sys


def main():
    path, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else '')
    ms, n = collections.Counter(), collections.Counter()
    for row in csv.reader(l for l in open(path) if l.startswith('"')):
        if row[0] == 'ID':
            continue
        name = re.sub(r'[<(].*', '', row[4].replace('<unnamed>::', '')).split('::')[-1]
        ms[name] += float(row[-1]) / 1e6
        n[name] += 1
    total = sum(ms.values())
    print(f'# ncu launch list of one training step{": " + title if title else ""}\n')
    print('`python bench.py --ncu` (exit 0), then the same command under `ncu --metrics gpu__time_duration.sum --clock-control none '
          '-s 2800 -c 2800`\n(cold cache, serialised: compare shares, not absolutes).  Raw list: `' + path.split('/')[-1] + '`.\n')
    print(f'{sum(n.values())} launches, {total:.2f} ms serialised.\n')
    print('| kernel | launches | ms | share |\n|---|---|---|---|')
    for k, v in ms.most_common():
        print(f'| `{k}` | {n[k]} | {v:.2f} | {100 * v / total:.1f} % |')


if __name__ == '__main__':
    main()
