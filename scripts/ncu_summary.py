"""Summarise an .ncu-rep: per kernel key metrics + opcode histogram + execution-count clusters."""
import csv, collections, subprocess, sys, io

def raw(rep):
	out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
	r = list(csv.reader(io.StringIO(out)))
	h = r[0]
	keys = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
		'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
		'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
		'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
		'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
		'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'l1tex__data_pipe_lsu_wavefronts.sum',
		'sm__inst_executed_pipe_lsu.sum', 'smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct',
		'smsp__warp_issue_stalled_barrier_per_warp_active.pct', 'smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct',
		'smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct', 'smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct',
		'smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct', 'smsp__warp_issue_stalled_wait_per_warp_active.pct',
		'smsp__warp_issue_stalled_not_selected_per_warp_active.pct']
	for row in r[2:]:
		for k in keys:
			if k in h:
				print("  %-80s %s %s" % (k, row[h.index(k)], r[1][h.index(k)]))
		print()

def sass(rep, pattern):
	out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pattern], capture_output=True, text=True).stdout
	rows = list(csv.reader(io.StringIO(out)))
	hi = [i for i, r in enumerate(rows) if 'Instructions Executed' in r][0]
	h = rows[hi]
	ie = h.index('Instructions Executed'); so = h.index('Source'); sm = h.index('# Samples')
	data = []
	for r in rows[hi + 1:]:
		if len(r) > ie:
			try: data.append((int(r[ie]), int(r[sm]), r[so].strip()))
			except ValueError: break
	tot = sum(d[0] for d in data)
	print(pattern, 'total warp-inst', tot, 'sass lines', len(data))
	ops = collections.Counter()
	for n, s, src in data:
		parts = src.split()
		op = parts[1] if parts[0].startswith('@') else parts[0]
		ops[op.split('.')[0]] += n
	print('  ' + ', '.join('%s %.1f%%' % (k, 100 * v / tot) for k, v in ops.most_common(24)))
	levels = collections.Counter()
	for n, s, src in data: levels[n] += 1
	print('  clusters (exec count x n_instr):', sorted(levels.items(), reverse=True)[:14])
	return data

if __name__ == "__main__":
	rep = sys.argv[1]
	raw(rep)
	for pat in sys.argv[2:]:
		sass(rep, pat)
