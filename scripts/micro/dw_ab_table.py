"""Table of a scripts/micro/dw_ab.sh log: python scripts/micro/dw_ab_table.py gpurun_out/dw_ab2.log"""
import re
import sys

txt = open(sys.argv[1]).read()
cur, res = None, {}
names = ['gru.dw7', 'convc1.dw15', 'conv.dw15', 'fhead.dw15']
for line in txt.splitlines():
    m = re.match(r'== variant (\S+) TH=(\d+) VFML_DW_TMA=(\d): (\w+)', line)
    if m:
        cur = m.groups()
        continue
    if cur and cur[3] == 'tests' and re.search(r'passed|failed|error', line, re.I):
        print('tests', cur[:3], line.strip())
    m = re.match(r'(\S+)\s+C=\s*\d+ n_img=\d\s+([\d.]+) us', line)
    if m and cur and cur[3] == 'timing':
        res.setdefault(cur[:3], {})[m.group(1)] = float(m.group(2))
    if cur and cur[3] == 'timing' and 'Error' in line:
        res.setdefault(cur[:3], {})['err'] = line.strip()[:60]
print('variant TH tma ' + ' '.join(f'{n:>12s}' for n in names) + '      sum')
for k, v in res.items():
    tot = sum(v.get(n, 0) for n in names)
    print(f'{k[0]:6s} {k[1]:>3s} {k[2]:>3s} ' + ' '.join(f'{v[n]:12.1f}' if n in v else f'{"-":>12s}' for n in names) + f' {tot:8.1f} ' + v.get('err', ''))
