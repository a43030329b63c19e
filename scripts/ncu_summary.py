#!/usr/bin/env python
"""Turn an Nsight Compute report into the markdown summary committed under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof_gemm.ncu-rep [--kernel REGEX] [--out profiles/ncu_gemm.md]

Reads the report with `ncu -i ... --page raw --csv` (works on the GPU-less build box), keeps one
row per profiled launch and the metrics the profiling recipe asks for (duration, launch shape,
registers, tensor-pipe / TMEM / DRAM / L2 / SM busy percentages, achieved occupancy), then the top
warp-stall reasons.  Nothing is measured here: timing numbers quoted elsewhere never come from
a run under the profiler.
"""
import argparse
import csv
import io
import re
import subprocess
import sys

# (column label, candidate metric names: first one present in the report wins)
WANTED = [
    ('duration us', ['gpu__time_duration.sum']),
    ('regs/thread', ['launch__registers_per_thread']),
    ('dyn smem KB', ['launch__shared_mem_per_block_dynamic']),
    ('SM busy %', ['sm__throughput.avg.pct_of_peak_sustained_elapsed']),
    ('tensor pipe %', ['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
                       'sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active']),
    ('TMEM %', ['sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active']),
    ('FMA pipe %', ['sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
                    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active']),
    ('DRAM %', ['gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
                'dram__throughput.avg.pct_of_peak_sustained_elapsed']),
    ('DRAM read', ['dram__bytes_read.sum']),
    ('DRAM write', ['dram__bytes_write.sum']),
    ('L2 %', ['lts__throughput.avg.pct_of_peak_sustained_elapsed']),
    ('L1/TEX %', ['l1tex__throughput.avg.pct_of_peak_sustained_elapsed']),
    ('occupancy %', ['sm__warps_active.avg.pct_of_peak_sustained_active']),
    ('IPC', ['sm__inst_executed.avg.per_cycle_active']),
]
STALL = re.compile(r'smsp__average_warps?_issue_stalled_(\w+?)_per_issue_active|smsp__average_warp_latency_issue_stalled_(\w+)')


def read_raw(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True)
    if out.returncode != 0:
        sys.exit('ncu failed: ' + out.stderr[:400])
    rows = list(csv.reader(io.StringIO(out.stdout)))
    head, units, data = rows[0], rows[1], rows[2:]
    return head, units, data


def col_index(head, name, data=None):
    """Column of metric ``name``; the same metric can appear under several sections, some of them
    not collected (empty cells): prefer a column that has a value."""
    hits = [i for i, h in enumerate(head) if h == name or h.endswith('.' + name)]
    if data:
        for i in hits:
            if any(len(row) > i and row[i] != '' for row in data):
                return i
    return hits[0] if hits else None


def fmt(v, unit):
    try:
        x = float(v.replace(',', ''))
    except (ValueError, AttributeError):
        return v or '-'
    if unit in ('ns', 'nsecond'):
        return '{:.1f}'.format(x / 1e3)
    if unit in ('us', 'usecond'):
        return '{:.1f}'.format(x)
    if unit in ('ms', 'msecond'):
        return '{:.1f}'.format(x * 1e3)
    if unit == 'byte':
        return '{:.1f} KB'.format(x / 1024) if x < 1e7 else '{:.1f} MB'.format(x / 1e6)
    if unit in ('Kbyte', 'Mbyte', 'Gbyte'):
        return '{:.1f} {}B'.format(x, unit[0])
    return '{:.1f}'.format(x) if abs(x) < 1e5 else '{:.3g}'.format(x)


def hot_instructions(path, kernel_regex, top):
    """Top SASS instructions by warp-stall samples of the first matching launch (source page)."""
    out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kernel_regex,
                          '--launch-count', '1'], capture_output=True, text=True)
    if out.returncode != 0:
        return []
    rows = list(csv.reader(io.StringIO(out.stdout)))
    hdr = next((i for i, r in enumerate(rows) if r and r[0] == 'Address'), None)
    if hdr is None:
        return []
    head = rows[hdr]
    try:
        i_src, i_smp = head.index('Source'), head.index('# Samples')
    except ValueError:
        return []
    inst, seen = [], set()
    for k, r in enumerate(rows[hdr + 1:]):
        if len(r) <= i_smp or r[0] == 'Address' or not r[0].startswith('0x') or r[0] in seen:
            continue
        seen.add(r[0])       # the page can list the same code twice (one block per view)
        try:
            inst.append((int(r[i_smp]), k, r[i_src].strip()))
        except ValueError:
            pass
    total = sum(n for n, _, _ in inst) or 1
    lines = ['', 'Hottest SASS instructions of the first matching launch (warp-stall samples, {} total):'.format(total), '',
             '| samples | share | instruction # | SASS |', '|---|---|---|---|']
    for n, k, src in sorted(inst, reverse=True)[:top]:
        lines.append('| {} | {:.1f} % | {} | `{}` |'.format(n, 100.0 * n / total, k, src.replace('|', '\\|')))
    return lines


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('report')
    ap.add_argument('--kernel', default='.', help='regex on the kernel name')
    ap.add_argument('--out', default='')
    ap.add_argument('--title', default='')
    ap.add_argument('--hot', type=int, default=12, help='how many hottest SASS instructions to list (0 = none)')
    a = ap.parse_args()
    head, units, data = read_raw(a.report)
    k_name, k_grid, k_block = col_index(head, 'Kernel Name'), col_index(head, 'Grid Size'), col_index(head, 'Block Size')
    picked = []
    for label, names in WANTED:
        for n in names:
            i = col_index(head, n, data)
            if i is not None:
                picked.append((label, i))
                break
    stall_cols = []
    for i, h in enumerate(head):
        m = STALL.search(h)
        if m and 'not_issued' not in h:
            stall_cols.append(((m.group(1) or m.group(2)), i))
    lines = ['# {}'.format(a.title or 'ncu summary of ' + a.report.split('/')[-1]), '',
             'Generated by `scripts/ncu_summary.py` from `{}` (`ncu --set full`, read with `--page raw --csv`). '
             'Durations are under the profiler (replay, clocks not locked) and are never quoted as benchmark values.'.format(
                 a.report.split('/')[-1]), '',
             '| # | kernel | grid x block | ' + ' | '.join(l for l, _ in picked) + ' |',
             '|---|---|---|' + '---|' * len(picked)]
    stalls_out = []
    for r, row in enumerate(data):
        if not row or not re.search(a.kernel, row[k_name]):
            continue
        name = re.sub(r'\(.*', '', row[k_name]).replace('<unnamed>::', '').replace('(anonymous namespace)::', '')
        cells = [fmt(row[i], units[i]) for _, i in picked]
        lines.append('| {} | `{}` | {} x {} | {} |'.format(r, name, row[k_grid], row[k_block], ' | '.join(cells)))
        st = []
        for reason, i in stall_cols:
            try:
                st.append((float(row[i]), reason))
            except ValueError:
                pass
        st.sort(reverse=True)
        tot = sum(v for v, _ in st) or 1.0
        stalls_out.append('* launch {}: '.format(r) + ', '.join('{} {:.0f} %'.format(n, 100 * v / tot) for v, n in st[:5]))
    lines += ['', 'Top warp-stall reasons (share of stalled issue slots):', ''] + stalls_out
    if a.hot > 0:
        lines += hot_instructions(a.report, a.kernel, a.hot)
    text = '\n'.join(lines) + '\n'
    if a.out:
        with open(a.out, 'w') as f:
            f.write(text)
    print(text)


if __name__ == '__main__':
    main()
