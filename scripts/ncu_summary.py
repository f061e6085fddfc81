"""Compact text summary of an .ncu-rep (--set full): one block per selected launch with the metrics the roofline
discussion uses.  Usage: ncu_summary.py report.ncu-rep [launch indices ...] > profiles/<name>.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
want = [int(a) for a in sys.argv[2:]]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
KEYS = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'lts__t_sectors_srcunit_tex_op_read.sum',
        'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor_subpipe_hmma.sum', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed']
print(f'# {rep}: {len(data)} launches captured with ncu --set full --clock-control none')
for i, r in enumerate(data):
    if want and i not in want:
        continue
    print(f'\n== launch {i}: {r[hdr.index("Kernel Name")][:90]}')
    for k in KEYS:
        if k in hdr:
            j = hdr.index(k)
            print(f'  {k:80s} {r[j]:>16s} {units[j]}')
