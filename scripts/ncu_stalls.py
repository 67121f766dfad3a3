"""Top stall locations (SASS) of one kernel in an .ncu-rep."""
import csv, collections, subprocess, io, sys

def top(rep, pat, n=40):
	out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
	rows = list(csv.reader(io.StringIO(out)))
	hi = [i for i, r in enumerate(rows) if 'Instructions Executed' in r][0]
	h = rows[hi]
	so = h.index('Source'); sm = h.index('# Samples'); ie = h.index('Instructions Executed')
	stall_cols = [i for i, k in enumerate(h) if k.startswith('stall_') and 'Not Issued' not in k]
	data = []
	for r in rows[hi + 1:]:
		if len(r) > ie and r[sm].isdigit():
			data.append(r)
		elif len(r) > ie:
			break
	tot = sum(int(r[sm]) for r in data)
	print(pat, 'total samples', tot)
	agg = collections.Counter()
	for r in data:
		for i in stall_cols:
			if r[i].isdigit(): agg[h[i]] += int(r[i])
	print('  stall totals:', ', '.join('%s %.1f%%' % (k, 100 * v / max(tot, 1)) for k, v in agg.most_common(10)))
	cl = collections.Counter()
	for r in data: cl[int(r[ie])] += int(r[sm])
	print('  samples by exec-count cluster:', [(k, '%.1f%%' % (100 * v / tot)) for k, v in cl.most_common(8)])
	idx = sorted(range(len(data)), key=lambda i: -int(data[i][sm]))[:n]
	for i in sorted(idx):
		r = data[i]
		st = sorted(((int(r[c]) if r[c].isdigit() else 0, h[c]) for c in stall_cols), reverse=True)[:2]
		print('  %5d %8s %5.1f%% %-64s %s' % (i, r[ie], 100 * int(r[sm]) / tot, r[so].strip()[:64], st))

if __name__ == "__main__":
	top(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)
