"""ncu_brief.py REPORT.ncu-rep OUT.txt "header line": the key metrics of every kernel of an ncu
--set full capture as text (the same list make_profiles.py writes for the bench step)."""
import csv, io, subprocess, sys

rep, out_path, header = sys.argv[1], sys.argv[2], sys.argv[3]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(out)))
h, units = r[0], r[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
	'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
	'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
	'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'sm__cycles_active.avg', 'sm__cycles_active.max', 'sm__cycles_elapsed.max',
	'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
	'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
	'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
	'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
	'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
lines = ["# " + header, ""]
for row in r[2:]:
	short = row[h.index('Kernel Name')].split('(')[0].replace('void ', '').replace('<unnamed>::', '')
	lines.append("## " + short)
	for k in keys:
		if k in h:
			lines.append("%-82s %s %s" % (k, row[h.index(k)], units[h.index(k)]))
	for i, k in enumerate(h):
		if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio'):
			try:
				if float(row[i]) >= 0.1:
					lines.append("%-82s %s" % (k, row[i]))
			except ValueError:
				pass
	lines.append("")
open(out_path, 'w').write("\n".join(lines))
print(out_path, len(r) - 2, "kernels")
