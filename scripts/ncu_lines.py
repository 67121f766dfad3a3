"""Per CUDA source line: warp instructions executed and stall samples, from an .ncu-rep (development aid).
usage: ncu_lines.py REP [kernel-regex] [top-n]"""
import csv, io, subprocess, sys, collections

rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else None
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]
if pat:
	cmd += ["--kernel-name", "regex:" + pat]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if 'Instructions Executed' in r][0]
h = rows[hi]
ie, sm, ln = h.index('Instructions Executed'), h.index('# Samples'), h.index('Line No')
src_col = h.index('Source')
stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith('stall_') and 'Not Issued' not in c]
per = collections.OrderedDict()
cur = None
for r in rows[hi + 1:]:
	if len(r) <= ie:
		continue
	if r[ln].strip():
		cur = (r[ln].strip(), r[src_col].strip())
		per.setdefault(cur, [0, 0, collections.Counter()])
		continue
	try:
		n, s = int(r[ie]), int(r[sm])
	except ValueError:
		continue
	if cur is None:
		continue
	per[cur][0] += n
	per[cur][1] += s
	for i, c in stall_cols:
		try:
			per[cur][2][c] += int(r[i])
		except ValueError:
			pass
tot_i = sum(v[0] for v in per.values())
tot_s = sum(v[1] for v in per.values())
print("total warp-inst %d, samples %d" % (tot_i, tot_s))
key = (lambda kv: -kv[1][0]) if (len(sys.argv) > 4 and sys.argv[4] == "inst") else (lambda kv: -kv[1][1])
for (line, text), (n, s, st) in sorted(per.items(), key=key)[:top]:
	main = ", ".join("%s %d" % (k[6:], v) for k, v in st.most_common(3))
	print("%5s  inst %5.1f%%  samples %5.1f%%  [%s]  %s" % (line, 100 * n / max(tot_i, 1), 100 * s / max(tot_s, 1), main, text[:90]))
