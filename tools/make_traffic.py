"""profiles/traffic.json from an ncu launch list with per-launch traffic (tools/_run_prof_r02.sh):

    python tools/make_traffic.py profiles/r02f_launches.csv [n_edges]

Takes ONE full-size force call out of the list - the launches between two consecutive `partition_keys_kernel` launches
(the first kernel of a step of the 100,000-atom cell) - and sums duration, DRAM bytes and L2 bytes (sectors x 32) per
kernel and per bench category.  bench.py reads `dram_bytes_per_step[category]` for `roofline.traffic`."""
import collections
import csv
import json
import os
import sys

CATS = [('conv_pair_fwd', 'conv_fwd'), ('conv_fwd', 'conv_fwd'), ('conv_pair_bwd', 'conv_bwd_edge'), ('conv_bwd', 'conv_bwd_edge'),
        ('irreps_linear', 'node_linear_gate'), ('linear_prepack', 'node_linear_gate'), ('gate_', 'node_linear_gate'),
        ('block_transpose', 'node_linear_gate'), ('embed_fwd_kernel', 'node_linear_gate'), ('rescale', 'node_linear_gate'),
        ('colsum', 'node_linear_gate'), ('force_virial', 'force_virial'),
        ('edge_embed', 'graph+edge_embed+csr'), ('neighbor_kernel', 'graph+edge_embed+csr'), ('grid_setup', 'graph+edge_embed+csr'),
        ('bin_atoms', 'graph+edge_embed+csr'), ('fill_bins', 'graph+edge_embed+csr'), ('exclusive_scan', 'graph+edge_embed+csr'),
        ('histogram', 'graph+edge_embed+csr'), ('scatter_by_key', 'graph+edge_embed+csr'), ('sort_rows', 'graph+edge_embed+csr'),
        ('partition_keys', 'graph+edge_embed+csr'), ('csr_', 'graph+edge_embed+csr')]


def category(name):
    for key, cat in CATS:
        if key in name:
            return cat
    return 'other'


def short(name):
    n = name.split('(')[0].strip()
    for pre in ('void ', 'b200::', 'pairk::', '<unnamed>::'):
        n = n.replace(pre, '')
    return n[-70:]


def main():
    path = sys.argv[1]
    rows = list(csv.reader(open(path)))
    h = next(r for r in rows if 'Kernel Name' in r)
    start = rows.index(h)
    ki, mi, vi, ui, ii = (h.index(k) for k in ('Kernel Name', 'Metric Name', 'Metric Value', 'Metric Unit', 'ID'))
    launches = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) <= vi or not r[ii].isdigit():
            continue
        v = float(r[vi].replace(',', ''))
        u = r[ui]
        if r[mi] == 'gpu__time_duration.sum':
            v = v / 1e6 if u in ('ns', 'nsecond') else v / 1e3 if u in ('us', 'usecond') else v * 1e3 if u in ('s', 'second') else v
        elif 'byte' in u:
            v *= {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1.0)
        launches.setdefault(int(r[ii]), {'name': r[ki]})[r[mi]] = v
    ids = list(launches)
    marks = [i for i in ids if 'partition_keys_kernel' in launches[i]['name']]
    if len(marks) < 2:
        raise SystemExit('need two full-size steps in the list (partition_keys_kernel launches): found %d' % len(marks))
    a, b = marks[-2], marks[-1]
    seg = [i for i in ids if a <= i < b]
    per = collections.OrderedDict()
    dram, l2, ms = collections.defaultdict(float), collections.defaultdict(float), collections.defaultdict(float)
    for i in seg:
        e = launches[i]
        c = category(e['name'])
        d = e.get('dram__bytes_read.sum', 0.0) + e.get('dram__bytes_write.sum', 0.0)
        s = 32.0 * e.get('lts__t_sectors.sum', 0.0)
        t = e.get('gpu__time_duration.sum', 0.0)
        dram[c] += d
        l2[c] += s
        ms[c] += t
        k = per.setdefault(short(e['name']), {'launches': 0, 'ms': 0.0, 'dram_bytes': 0.0, 'l2_bytes': 0.0})
        k['launches'] += 1
        k['ms'] += t
        k['dram_bytes'] += d
        k['l2_bytes'] += s
    out = {'n_edges': int(sys.argv[2]) if len(sys.argv) > 2 else 4808858,
           'dram_bytes_per_step': dict(dram), 'l2_bytes_per_step': dict(l2), 'ms_per_step_under_ncu': dict(ms),
           'source': f'{path}: ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum (x 32 B) '
                     f'--clock-control none over python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune --no-alt; summed '
                     f'over the {len(seg)} launches of ONE 100,000-atom force call (launches {a}-{b - 1})',
           'per_kernel': per}
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    json.dump(out, open(os.path.join(root, 'profiles', 'traffic.json'), 'w'), indent=1)
    print(len(seg), 'launches,', round(sum(ms.values()), 2), 'ms')
    for c in ms:
        print('%-24s %8.2f ms  %8.2f GB DRAM  %8.2f GB L2' % (c, ms[c], dram[c] / 1e9, l2[c] / 1e9))


if __name__ == '__main__':
    main()
