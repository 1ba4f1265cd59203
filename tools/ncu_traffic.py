"""profiles/ncu_traffic.json from the raw page of an `ncu --set full` capture: DRAM bytes (read + write) per launch of the
kernels named on the command line, stamped with the commit, kernel name and workload the capture was taken on, so that
bench.py can refuse a figure that belongs to another kernel.

    python tools/ncu_traffic.py gpurun_out/<tag>/ncu_full_raw.csv C2/N16384 [kernel substring ...]
"""
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCALE = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}


def main():
    raw, workload = sys.argv[1], sys.argv[2]
    wanted = sys.argv[3:] or ['gp_trisumsq_ozaki', 'gp_kstar_slices8']
    rows = list(csv.reader(open(raw)))
    head, units = rows[0], rows[1]
    name, rd, wr = head.index('Kernel Name'), head.index('dram__bytes_read.sum'), head.index('dram__bytes_write.sum')
    dur = head.index('gpu__time_duration.sum')
    commit = subprocess.run(['git', '-C', ROOT, 'rev-parse', '--short=12', 'HEAD'], capture_output=True,
                            text=True).stdout.strip()
    per_kernel = {}
    for r in rows[2:]:
        for sub in wanted:
            if sub in r[name]:
                # "se::oz::gp_trisumsq_ozaki_kernel<6, 2>(args...)" -> "oz::gp_trisumsq_ozaki_kernel<6,2>"
                kname = re.sub(r'\(.*$', '', r[name]).replace('se::', '', 1).replace(', ', ',').replace('void ', '')
                if '<' in kname and 'kstar' in kname:
                    kname = kname.split('<')[0]
                if not kname.startswith('oz::'):  # the raw page prints the bare function name; bench.py asks with the namespace
                    kname = 'oz::' + kname
                e = per_kernel.setdefault(kname, dict(bytes=[], us=[]))
                e['bytes'].append(float(r[rd]) * SCALE[units[rd]] + float(r[wr]) * SCALE[units[wr]])
                e['us'].append(float(r[dur]) * {'ns': 1e-3, 'us': 1.0, 'ms': 1e3}.get(units[dur], 1.0))
    out = [dict(kernel=k, workload=workload, commit=commit, source=os.path.relpath(raw, ROOT),
                dram_bytes_per_launch=sum(v['bytes']) / len(v['bytes']), launches_captured=len(v['bytes']),
                duration_us_under_ncu=sum(v['us']) / len(v['us'])) for k, v in per_kernel.items()]
    path = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    old = []
    if os.path.exists(path):
        old = [e for e in json.load(open(path)) if (e['kernel'], e['workload']) not in {(o['kernel'], o['workload']) for o in out}]
    json.dump(old + out, open(path, 'w'), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
