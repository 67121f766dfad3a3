"""Screened (fp32 + float64 fallback) against exact float64 strip extraction: equality on
ordinary and adversarial dense maps, and timing (development aid).  The BBOX variant of the
kernel always takes the exact path, so both run from the same library."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from origami_b200 import ops, synth
from probe_perf import timeit

def main():
	W, H, cell, n_lines = (int(os.environ.get(k, d)) for k, d in (('PW', 4800), ('PH', 7200), ('PCELL', 25), ('PLINES', 2000)))
	dev = torch.device("cuda:0")
	page = synth.make_page(W, H, seed=0, kind="text")
	grid = synth.make_grid(W, H, cell, seed=0)
	gh, gw = grid.shape[:2]
	lines = synth.make_lines(gw * cell, gh * cell, n_lines, seed=0, dpi=600)
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	mp = ops.expand_map(g, cell)
	la = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines])
	ld = ops.lines_to_device(la, dev)
	plan = ops.line_steps(g, cell, ld, n_lines)
	total = int(plan.pix_offsets[-1].item())
	s1 = torch.zeros((total,), dtype=torch.uint8, device=dev)
	s2 = torch.zeros((total,), dtype=torch.uint8, device=dev)
	bbox = torch.empty((n_lines, 4), dtype=torch.int32, device=dev)
	gen = torch.Generator(device=dev); gen.manual_seed(5)
	Hd, Wd = mp.shape[:2]
	maps = {"synthetic": mp}
	maps["noise 0.5px"] = mp + 0.5 * torch.randn(mp.shape, device=dev, generator=gen)
	maps["noise 30px"] = mp + 30 * torch.randn(mp.shape, device=dev, generator=gen)
	maps["shift -40"] = mp - 40.0
	maps["x1.7 (partly outside)"] = mp * 1.7
	maps["constant"] = torch.full_like(mp, 1234.5)
	maps["half-integers/32"] = torch.round(mp * 64) / 64
	maps["tiny"] = mp * 1e-3
	maps["huge"] = mp * 1e6
	nanmap = mp.clone(); nanmap[::97, ::89] = float("nan"); maps["nan holes"] = nanmap
	infmap = mp.clone(); infmap[::101, ::83] = float("inf"); infmap[::103, ::79] = -float("inf"); maps["inf holes"] = infmap
	maps["around 4096 (ulp step)"] = 4096.0 + (mp - mp.mean()) * 0.01
	maps["around 8192 (off page)"] = 8192.0 + (mp - mp.mean()) * 0.003
	maps["steep x50"] = (mp - mp.mean()) * 50.0 + 2000.0
	maps["nearly flat"] = 1000.0 + mp * 1e-6
	maps["page edges"] = torch.stack([mp[..., 0] * ((W - 1) / float(mp[..., 0].max())), mp[..., 1] * ((H - 1) / float(mp[..., 1].max()))], dim=-1)
	maps["ties 1/32 + noise"] = torch.round(mp * 32) / 32 + 1.0 / 64 + 1e-6 * torch.randn(mp.shape, device=dev, generator=gen)
	maps["negative zero"] = mp * -0.0
	if "--kernel-only" in sys.argv:
		for _ in range(3):
			ops.extract_lines(src, g, cell, plan, s1, dense_map=mp)
		torch.cuda.synchronize()
		return
	if "--time-only" in sys.argv:
		maps = {}
	for name, m in maps.items():
		m = m.contiguous()
		s1.zero_(); s2.zero_()
		ops.extract_lines(src, g, cell, plan, s1, dense_map=m)
		ops.extract_lines(src, g, cell, plan, s2, bbox, dense_map=m)
		torch.cuda.synchronize()
		diff = int((s1 != s2).sum().item())
		print("%-24s mismatches %d of %d" % (name, diff, total))
	for name, fn in [
		("extract map screened", lambda: ops.extract_lines(src, g, cell, plan, s1, dense_map=mp)),
		("extract map exact+bbox", lambda: ops.extract_lines(src, g, cell, plan, s2, bbox, dense_map=mp)),
	]:
		med, best = timeit(fn)
		print("%-24s median %.3f ms best %.3f ms  (%.2f ns per 1000 samples, %d samples)" % (name, med, best, med * 1e9 / total, total))

if __name__ == "__main__":
	main()
