"""
ORACLE (test infrastructure).  Generates tests/golden/*.npz by running the UNMODIFIED
reference (imported from /root/reference through oracle/ref_import.py) on small seeded
synthetic artifacts.  Run in the build container only:

    python -m oracle.make_golden

The fp32 dense map is produced with cv2.setUseOptimized(False) (OpenCV's generic code
path, deterministic operation order); the IPP-dispatched default path differs from it by a
few ulp and its deviation is stored as `map_ipp_maxdiff` for information.
"""

import os
import sys

import cv2
import numpy as np
import PIL.Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_import  # noqa: E402
from origami_b200 import synth  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def make_case(name, width, height, cell, n_lines, seed, kind):
	Grid, Dewarper, Line = ref_import.load()
	page = synth.make_page(width, height, seed=seed, kind=kind, dpi=300)
	grid_hv = synth.make_grid(width, height, cell, seed=seed)

	cv2.setUseOptimized(True)
	ipp_map = Grid(grid_hv.copy(), cell).points("full")
	cv2.setUseOptimized(False)
	grid = Grid(grid_hv.copy(), cell)
	pil = PIL.Image.fromarray(page)
	dewarper = Dewarper(pil, grid)
	dense = grid.points("full").copy()
	dewarped = np.array(dewarper.dewarped)

	labels = (synth.make_labels(max(width // 2, 8), max(height // 2, 8), seed=seed)).astype(np.uint8)
	# fresh Grid: dewarp_image scales the cached map in place (reference bug)
	d2 = Dewarper(pil, Grid(grid_hv.copy(), cell))
	labels_nearest = np.array(d2.dewarp_image(PIL.Image.fromarray(labels * 60), cv2.INTER_NEAREST))
	d3 = Dewarper(pil, Grid(grid_hv.copy(), cell))
	labels_linear = np.array(d3.dewarp_image(PIL.Image.fromarray(labels * 60), cv2.INTER_LINEAR))

	Hd, Wd = dense.shape[:2]
	lines = synth.make_lines(Wd, Hd, n_lines, seed=seed, dpi=150)
	block = ref_import._Block(ref_import._Page(pil, dewarper))
	out = dict(page=page, grid=grid_hv, cell=np.int32(cell), dense=dense, dewarped=dewarped,
		labels=labels, labels_nearest=labels_nearest, labels_linear=labels_linear,
		map_ipp_maxdiff=np.float32(np.abs(ipp_map - dense).max()),
		n_lines=np.int32(len(lines)))
	rng = np.random.default_rng(seed)
	pts = np.stack([rng.uniform(-15, Hd + 15, 400), rng.uniform(-15, Wd + 15, 400)], axis=1)
	pts[:40] = np.rint(pts[:40])
	out["inv_pts"] = pts
	out["inv_out"] = grid.inverse_yx(pts)
	for i, ln in enumerate(lines):
		L = ref_import.make_line(Line, block, ln["p"], ln["right"], ln["up"])
		out["line%d_pru" % i] = np.array([ln["p"], ln["right"], ln["up"]], dtype=np.float64)
		out["line%d_strip48" % i] = np.array(L.dewarped_image(48))
		out["line%d_strip11" % i] = np.array(L.dewarped_image(11))
		out["line%d_half" % i] = L.warped_grid(xres=0.5, yres=0.5)
		path, height = L.warped_path(1)
		out["line%d_path" % i] = path
		out["line%d_pathh" % i] = np.float64(height)
		# ConfidenceSampler arithmetic (batch/detect/lines.py:40-44) on the reference grid
		g = L.warped_grid(xres=0.5, yres=0.5)
		g[:, :, 0] *= labels.shape[1] / page.shape[1]
		g[:, :, 1] *= labels.shape[0] / page.shape[0]
		lab = cv2.remap(labels, g, None, cv2.INTER_NEAREST)
		out["line%d_counts" % i] = np.bincount(lab.flatten(), minlength=4)
	cv2.setUseOptimized(True)
	os.makedirs(GOLDEN, exist_ok=True)
	np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **out)
	print(name, "dense", dense.shape, "lines", len(lines), "ipp maxdiff", out["map_ipp_maxdiff"])


def main():
	if not ref_import.available():
		raise SystemExit("/root/reference is not present: golden vectors can only be regenerated in the build container")
	make_case("ref_noise_160x200_c10", 160, 200, 10, 10, seed=21, kind="noise")
	make_case("ref_text_250x175_c25", 250, 175, 25, 8, seed=22, kind="text")
	make_case("ref_noise_97x131_c7", 97, 131, 7, 6, seed=23, kind="noise")


if __name__ == "__main__":
	main()
