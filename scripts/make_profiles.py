"""Turn the ncu captures in gpurun_out/ into the committed summaries under profiles/.
usage: python scripts/make_profiles.py <round-tag>   (expects gpurun_out/prof_<tag>_final.ncu-rep,
gpurun_out/launches_<tag>.csv, gpurun_out/bench_<tag>.json)"""
import collections, csv, io, json, subprocess, sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
rep = "gpurun_out/prof_%s_final.ncu-rep" % tag
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(out)))
h, units = r[0], r[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
	'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
	'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
	'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
	'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
	'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
	'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
	'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
	'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio']
lines = ["# ncu --set full --clock-control none --import-source on (B200, 4800x7200 page, 2000 lines), tag %s" % tag,
	"# command: python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --page-sets 1   (one launch of each kernel)", ""]
traffic = {}
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
names = {'og_page_pipelined_kernel': 'og_page_kernel', 'og_page_kernel': 'og_page_kernel', 'og_strips_screened_kernel': 'og_strips_screened_kernel', 'og_lines_map_kernel': 'og_lines_map_kernel', 'og_otsu_strips_kernel': 'og_binarize',
	'og_line_steps_kernel': 'og_line_steps_kernel', 'og_line_scan_kernel': 'og_line_scan_kernel', 'og_line_plan_kernel': 'og_line_plan_kernel'}
for row in r[2:]:
	short = row[h.index('Kernel Name')].split('(')[0].replace('void ', '').replace('<unnamed>::', '')
	lines.append("## " + short)
	for k in keys:
		if k in h:
			lines.append("%-82s %s %s" % (k, row[h.index(k)], units[h.index(k)]))
	rd = float(row[h.index('dram__bytes_read.sum')]) * scale[units[h.index('dram__bytes_read.sum')]]
	wr = float(row[h.index('dram__bytes_write.sum')]) * scale[units[h.index('dram__bytes_write.sum')]]
	lines.append("dram traffic (read+write)                                                          %.1f MB" % ((rd + wr) / 1e6))
	lines.append("")
	for a, b in names.items():
		if a in short:
			traffic[b] = int(rd + wr)
open('profiles/%s_ncu_full_summary.txt' % tag, 'w').write("\n".join(lines))
json.dump(traffic, open('profiles/traffic.json', 'w'), indent=1)

rows = [x for x in csv.reader(open('gpurun_out/launches_%s.csv' % tag)) if x and not x[0].startswith('==')]
hh = rows[0]
ki, vi, ii = hh.index('Kernel Name'), hh.index('Metric Value'), hh.index('ID')
agg = collections.defaultdict(list)
with open('profiles/%s_launches.csv' % tag, 'w') as f:
	f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare shares)\n")
	f.write("# command: python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --page-sets 1\n")
	f.write("id,kernel,duration_ns\n")
	for x in rows[1:]:
		nm = x[ki].split('(')[0].replace('void ', '').replace('<unnamed>::', '')
		f.write("%s,%s,%s\n" % (x[ii], nm, x[vi]))
		agg[nm].append(float(x[vi]))
tot = sum(sum(v) / len(v) for k, v in agg.items() if k.startswith('og_'))
for k, v in agg.items():
	if k.startswith('og_'):
		print("%-40s n=%d mean %.1f us share %.1f%%" % (k, len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / len(v) / tot))
bench = json.load(open('gpurun_out/bench_%s.json' % tag))
json.dump(bench, open('profiles/%s_bench_default.json' % tag, 'w'))
print("value", bench["value"], "ms/step", bench["ms_per_step"], "e2e", bench["e2e"]["value"])
