"""ncu report (--set full, one or more launches) -> a compact text digest for profiles/: one line per launch and one
block of averages per kernel.  usage: python tools/profile_digest.py report.ncu-rep > profiles/<name>.txt"""
import collections
import csv
import subprocess
import sys

COLS = [('gpu__time_duration.sum', 'us', 1.0), ('launch__grid_size', 'grid', 1.0), ('launch__block_size', 'block', 1.0),
        ('launch__registers_per_thread', 'regs', 1.0), ('launch__occupancy_limit_shared_mem', 'CTA/SM(smem)', 1.0),
        ('dram__bytes_read.sum', 'dram_rd', 1.0), ('dram__bytes_write.sum', 'dram_wr', 1.0),
        ('lts__t_bytes.sum', 'l2_bytes', 1.0),
        ('sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active', 'hmma%', 1.0),
        ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue%', 1.0),
        ('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'alu%', 1.0),
        ('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'fma%', 1.0),
        ('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'lsu%', 1.0),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps%', 1.0),
        ('dram__throughput.avg.pct_of_peak_sustained_elapsed', 'dram%', 1.0),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smem_conflicts', 1.0),
        ('smsp__inst_executed.sum', 'warp_inst', 1.0)]


def main():
    rep = sys.argv[1]
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = [(k, n) for k, n, _ in COLS if k in ix]
    print(f'# {rep}: {len(data)} launches; ncu --set full --clock-control none (cold caches, serialised: compare shares)')
    print('# units: ' + ', '.join(f'{n}[{units[ix[k]]}]' for k, n in cols))
    per = collections.OrderedDict()
    for r in data:
        name = r[ix['Kernel Name']].split('(')[0].replace('void ', '').replace('geotx::', '')
        vals = [r[ix[k]] for k, _ in cols]
        print(f'{name:42s} ' + ' '.join(f'{n}={v}' for (k, n), v in zip(cols, vals)))
        per.setdefault(name, []).append(float(r[ix['gpu__time_duration.sum']].replace(',', '')))
    total = sum(sum(v) for v in per.values())
    print('\n# per kernel: launches, mean us, total us, share of the serialised forward')
    for name, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
        print(f'{name:42s} x{len(v):3d}  {sum(v) / len(v):8.1f}  {sum(v):8.1f}  {100 * sum(v) / total:5.1f} %')
    print(f'{"total":42s} x{sum(len(v) for v in per.values()):3d}  {"":8s}  {total:8.1f}')


main()
