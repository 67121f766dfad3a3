"""
ORACLE (test infrastructure, not product code) -- bit-exact NumPy restatement of the
arithmetic on Origami's page-geometry hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  Nothing under origami_b200/ does.

Every function restates one row of SURVEY.md section 8(a) and cites the reference
call site (paths relative to /root/reference) whose behaviour it reproduces.  The
arithmetic that lives in third-party libraries the reference calls is restated from
the published algorithm:

  * OpenCV (reference pins opencv-python==4.2.0.32, docker/pip_requirements.txt;
    this image has 4.13.0): cv2.resize INTER_LINEAR generic path
    (modules/imgproc/src/resize.cpp: HResizeLinear / VResizeLinear), cv2.remap
    INTER_LINEAR / INTER_NEAREST on 8-bit images (modules/imgproc/src/imgwarp.cpp:
    RemapInvoker, remapBilinear, remapNearest), cv2.warpAffine.
  * SciPy RegularGridInterpolator, method="linear" (reference pins scipy==1.3.1; this
    image has 1.18.1): scipy/interpolate/_rgi.py _evaluate_linear.
  * scikit-image 0.16.2 threshold_sauvola / threshold_otsu (skimage/filters/
    thresholding.py); scikit-image is NOT installed here, so these two restatements
    are "parity unpinned" against skimage itself (they are pinned against a second,
    independent integral-image formulation in tests/test_oracle.py).

Pinning: tests/test_oracle.py checks every function here against cv2 (generic path,
cv2.setUseOptimized(False)), scipy and numpy run in-process, and against the golden
vectors in tests/golden/ that oracle/make_golden.py produced by importing the
UNMODIFIED reference modules from /root/reference.
"""

import math

import numpy as np

F32 = np.float32
F64 = np.float64

INTER_NEAREST = 0  # same integer values as cv2.INTER_NEAREST / cv2.INTER_LINEAR
INTER_LINEAR = 1


# --------------------------------------------------------------------------------------
# A2  Grid.points("full")        origami/core/dewarp.py:514-526
# --------------------------------------------------------------------------------------

def _resize_axis(n_src, s):
	"""Per destination index: source index and fp32 fraction, cv2.resize INTER_LINEAR.

	resize.cpp: fx = (float)((dx+0.5)*scale_x - 0.5) with scale_x = 1./(dsize/ssize)
	evaluated in double; sx = cvFloor(fx); fx -= sx (float).
	"""
	d = np.arange(n_src * s, dtype=F64)
	f = ((d + 0.5) * (1.0 / float(s)) - 0.5).astype(F32)
	i = np.floor(f).astype(np.int32)
	fr = (f - i.astype(F32)).astype(F32)
	return i, fr


def expand_plane(S, s):
	"""cv2.resize(S, (gw*s, gh*s), INTER_LINEAR) of one fp32 plane, generic code path.

	Horizontal pass first (x index clamped AND fraction reset at both borders), then
	vertical pass (row indices clamped, fraction NOT reset); every blend is
	fl(fl(a*(1-f)) + fl(b*f)) in fp32 without fused multiply-add.
	"""
	S = np.ascontiguousarray(S, dtype=F32)
	gh, gw = S.shape
	ix, fx = _resize_axis(gw, s)
	lo = ix < 0
	ix[lo] = 0
	fx[lo] = 0
	hi = ix >= gw - 1
	ix[hi] = gw - 1
	fx[hi] = 0
	ix1 = np.minimum(ix + 1, gw - 1)
	a0 = (F32(1) - fx).astype(F32)
	H = (S[:, ix] * a0).astype(F32) + (S[:, ix1] * fx).astype(F32)

	iy, fy = _resize_axis(gh, s)
	r0 = np.clip(iy, 0, gh - 1)
	r1 = np.clip(iy + 1, 0, gh - 1)
	b0 = (F32(1) - fy).astype(F32)[:, None]
	b1 = fy[:, None]
	return ((H[r0] * b0).astype(F32) + (H[r1] * b1).astype(F32)).astype(F32)


def expand_map(grid, s):
	"""Grid.points("full"): (gh, gw, 2) fp32 -> (gh*s, gw*s, 2) fp32, [..., 0] = x.

	origami/core/dewarp.py:518-524 resizes the x plane and the y plane separately and
	stacks them with np.dstack.
	"""
	grid = np.asarray(grid, dtype=F32)
	return np.dstack([expand_plane(grid[:, :, 0], s), expand_plane(grid[:, :, 1], s)])


# --------------------------------------------------------------------------------------
# A3 / A4  cv2.remap on 8-bit images   origami/core/dewarp.py:608-627, core/block.py:256
# --------------------------------------------------------------------------------------

def _sat_short(v):
	return np.clip(v, -32768, 32767)


def _tap(src, iy, ix, border):
	H, W = src.shape
	inside = (ix >= 0) & (ix < W) & (iy >= 0) & (iy < H)
	v = src[np.clip(iy, 0, H - 1), np.clip(ix, 0, W - 1)].astype(np.int64)
	return np.where(inside, v, border)


def remap_linear(src, map_xy, border=0):
	"""cv2.remap(src u8, map fp32 (h,w,2), None, INTER_LINEAR), BORDER_CONSTANT.

	imgwarp.cpp RemapInvoker: sx = cvRound(x*32) (fp32 product, round-half-even),
	ix = saturate_cast<short>(sx >> 5), fx = sx & 31; remapBilinear with the 15-bit
	BilinearTab_i weights, which for a 32x32 table are exactly (32-fx)(32-fy)*32 etc.,
	so out = (sum + 2^14) >> 15 == (sum/32 + 512) >> 10.  Taps outside the image
	contribute the border value.
	"""
	src = np.asarray(src, dtype=np.uint8)
	m = np.asarray(map_xy, dtype=F32)
	sx = np.rint(m[..., 0] * F32(32)).astype(np.int64)
	sy = np.rint(m[..., 1] * F32(32)).astype(np.int64)
	ix = _sat_short(sx >> 5)
	iy = _sat_short(sy >> 5)
	fx = sx & 31
	fy = sy & 31
	p00 = _tap(src, iy, ix, border)
	p01 = _tap(src, iy, ix + 1, border)
	p10 = _tap(src, iy + 1, ix, border)
	p11 = _tap(src, iy + 1, ix + 1, border)
	acc = ((32 - fx) * (32 - fy) * p00 + fx * (32 - fy) * p01
		+ (32 - fx) * fy * p10 + fx * fy * p11)
	return ((acc + 512) >> 10).astype(np.uint8)


def remap_nearest(src, map_xy, border=0):
	"""cv2.remap(..., INTER_NEAREST): p[rint(y), rint(x)], round-half-even, OOB -> border.

	imgwarp.cpp RemapInvoker nearest branch: XY = saturate_cast<short>(float).
	"""
	src = np.asarray(src, dtype=np.uint8)
	m = np.asarray(map_xy, dtype=F32)
	ix = _sat_short(np.rint(m[..., 0]).astype(np.int64))
	iy = _sat_short(np.rint(m[..., 1]).astype(np.int64))
	return _tap(src, iy, ix, border).astype(np.uint8)


def scale_map(map_xy, w0, h0, w1, h1):
	"""In-place-style scaling of an fp32 map by Python-float ratios.

	Dewarper.dewarp_image (origami/core/dewarp.py:622-623) and scale_grid
	(origami/batch/detect/lines.py:18-22): `grid[:, :, 0] *= w1 / w0` on an fp32 array
	multiplies in fp32 by the ratio rounded to fp32 (NumPy >= 2 weak-scalar rule).
	Returns a copy (the reference mutates the lru_cache'd map; that bug is not kept).
	"""
	m = np.array(map_xy, dtype=F32, copy=True)
	m[..., 0] *= F32(w1 / w0)
	m[..., 1] *= F32(h1 / h0)
	return m


def dewarp_page(src, grid, s):
	"""Dewarper.dewarped (origami/core/dewarp.py:608-614)."""
	return remap_linear(src, expand_map(grid, s))


def dewarp_image(im, page_size, grid, s, interpolation=INTER_LINEAR):
	"""Dewarper.dewarp_image (origami/core/dewarp.py:616-627); page_size = (w0, h0)."""
	w0, h0 = page_size
	h1, w1 = im.shape
	m = scale_map(expand_map(grid, s), w0, h0, w1, h1)
	if interpolation == INTER_NEAREST:
		return remap_nearest(im, m)
	return remap_linear(im, m)


# --------------------------------------------------------------------------------------
# A5  Grid.inverse_yx            origami/core/dewarp.py:561-567
# --------------------------------------------------------------------------------------

def inverse_yx(dense, pts_yx):
	"""RegularGridInterpolator((arange(Hd), arange(Wd)), dense, "linear",
	bounds_error=False, fill_value=None)(pts) -- linear extrapolation outside.

	scipy/interpolate/_rgi.py: find_indices gives i = clip(searchsorted-1, 0, n-2) and
	the normalised distance y - i (cell width 1); _evaluate_linear accumulates, in
	float64 and in this order, v00*((1-a)(1-b)) + v01*((1-a)b) + v10*(a(1-b)) + v11*(ab),
	starting from 0.0, with the fp32 nodes promoted to float64.

	The cell of an exactly integral coordinate k is k-1 with a = 1 in SciPy and k with
	a = 0 here; both give the same float64 result because the vanishing terms are exact
	zeros.
	"""
	dense = np.asarray(dense)
	Hd, Wd = dense.shape[:2]
	p = np.asarray(pts_yx, dtype=F64)
	y = p[..., 0]
	x = p[..., 1]
	i0 = np.clip(np.floor(y), 0, Hd - 2).astype(np.int64)
	i1 = np.clip(np.floor(x), 0, Wd - 2).astype(np.int64)
	a = y - i0
	b = x - i1
	wa = 1 - a
	wb = 1 - b
	out = np.zeros(p.shape[:-1] + (dense.shape[2],), dtype=F64)
	for (di, dj, w) in ((0, 0, wa * wb), (0, 1, wa * b), (1, 0, a * wb), (1, 1, a * b)):
		out = out + dense[i0 + di, i1 + dj].astype(F64) * w[..., None]
	return out


# --------------------------------------------------------------------------------------
# A7 / A8  Line._position, Line.warped_grid     origami/core/block.py:195-248
# --------------------------------------------------------------------------------------

def _segment_intersection_dd(p1, p2, q1, q2):
	"""Intersection point of segments p1-p2 and q1-q2, or None when they do not meet in exactly one
	point -- the logic of GEOS's LineIntersector as shapely's `a.intersection(b)` reaches it
	(geos/algorithm/LineIntersector.cpp computeIntersect + CGAlgorithmsDD::intersection):
	orientation tests, an endpoint that lies on the other segment is returned with its own
	coordinates, a proper crossing is evaluated in homogeneous (determinant) form in extended
	precision and rounded to double.  GEOS uses double-double (~106 bits); decimal arithmetic with
	60 digits stands in for it here."""
	import decimal
	ctx = decimal.Context(prec=60)
	D = lambda v: decimal.Decimal(float(v))  # exact conversion of a double

	def orient(a, b, c):
		v = ctx.subtract(ctx.multiply(ctx.subtract(D(b[0]), D(a[0])), ctx.subtract(D(c[1]), D(a[1]))),
			ctx.multiply(ctx.subtract(D(b[1]), D(a[1])), ctx.subtract(D(c[0]), D(a[0]))))
		return (v > 0) - (v < 0)

	def env_overlap():
		return (max(min(p1[0], p2[0]), min(q1[0], q2[0])) <= min(max(p1[0], p2[0]), max(q1[0], q2[0]))
			and max(min(p1[1], p2[1]), min(q1[1], q2[1])) <= min(max(p1[1], p2[1]), max(q1[1], q2[1])))

	if not env_overlap():
		return None
	Pq1, Pq2 = orient(p1, p2, q1), orient(p1, p2, q2)
	if (Pq1 > 0 and Pq2 > 0) or (Pq1 < 0 and Pq2 < 0):
		return None
	Qp1, Qp2 = orient(q1, q2, p1), orient(q1, q2, p2)
	if (Qp1 > 0 and Qp2 > 0) or (Qp1 < 0 and Qp2 < 0):
		return None
	if Pq1 == 0 and Pq2 == 0 and Qp1 == 0 and Qp2 == 0:
		return None  # collinear: a segment or nothing, not a Point (a single shared end point is not a case here)
	if Pq1 == 0 or Pq2 == 0 or Qp1 == 0 or Qp2 == 0:
		if Pq1 == 0:
			return np.array(q1, dtype=F64)
		if Pq2 == 0:
			return np.array(q2, dtype=F64)
		if Qp1 == 0:
			return np.array(p1, dtype=F64)
		return np.array(p2, dtype=F64)
	px = ctx.subtract(D(p1[1]), D(p2[1]))
	py = ctx.subtract(D(p2[0]), D(p1[0]))
	pw = ctx.subtract(ctx.multiply(D(p1[0]), D(p2[1])), ctx.multiply(D(p2[0]), D(p1[1])))
	qx = ctx.subtract(D(q1[1]), D(q2[1]))
	qy = ctx.subtract(D(q2[0]), D(q1[0]))
	qw = ctx.subtract(ctx.multiply(D(q1[0]), D(q2[1])), ctx.multiply(D(q2[0]), D(q1[1])))
	x = ctx.subtract(ctx.multiply(py, qw), ctx.multiply(qy, pw))
	y = ctx.subtract(ctx.multiply(qx, pw), ctx.multiply(px, qw))
	w = ctx.subtract(ctx.multiply(px, qy), ctx.multiply(qx, py))
	return np.array([float(ctx.divide(x, w)), float(ctx.divide(y, w))], dtype=F64)


def line_position(p, right, up, xres=1, column=None):
	"""Line._position (origami/core/block.py:195-221) without shapely.

	The reference intersects the baseline with the two vertical segments x = x0, x = x1 (from
	py0 - 1 to py1 + 1) via shapely / GEOS, absent here: parity of the intersection point's last
	bits is unpinned.  Restated with GEOS's own formulation (_segment_intersection_dd), which is
	independent of the parametric form the product code uses; a miss keeps the endpoint, as
	`intersect_segments(..., default=...)` does.
	"""
	p0 = np.asarray(p, dtype=F64)
	right = np.asarray(right, dtype=F64)
	up = np.asarray(up, dtype=F64)
	if column is not None:
		p1 = p0 + right
		px0, px1 = min(p0[0], p1[0]), max(p0[0], p1[0])
		py0, py1 = min(p0[1], p1[1]), max(p0[1], p1[1])
		x0, x1 = column
		if x0 is None:
			x0 = px0
		if x1 is None:
			x1 = px1

		def hit(xv, default):
			c = _segment_intersection_dd(p0, p1, (xv, py0 - 1), (xv, py1 + 1))
			return default if c is None else c

		q0 = hit(x0, p0)
		q1 = hit(x1, p1)
		right = q1 - q0
		p0 = q0
		xres = xres * ((x1 - x0) / (px1 - px0))
	return p0, right, up, xres


def _linspace2(stop, num):
	"""np.linspace([0, 0], stop, num) restated (numpy/_core/function_base.py).

	step = stop/div; if ANY component's step is zero every component uses (i/div)*stop,
	else i*step; the last sample is forced to `stop`.
	"""
	stop = np.asarray(stop, dtype=F64)
	i = np.arange(num, dtype=F64)[:, None]
	div = num - 1
	if div > 0:
		step = stop / div
		if (step == 0).any():
			y = (i / div) * stop
		else:
			y = i * step
	else:
		y = i * stop
	y = y + 0.0
	if num > 1:
		y[-1] = stop
	return y


def _rough_extent(dense, p, right, up):
	"""Line.warped_grid(xsteps=2, ysteps=2) on the un-clipped line (block.py:226-228)."""
	g = lattice_to_warped(dense, np.asarray(p, F64), np.asarray(right, F64), np.asarray(up, F64), 2, 2)
	wx = np.max(np.abs(g[:, 0, 0] - g[:, 1, 0]))  # fp32
	wy = np.max(np.abs(g[0, :, 1] - g[1, :, 1]))  # fp32
	return wx, wy


def lattice_to_warped(dense, p0, right, up, xsteps, ysteps):
	"""The body of Line.warped_grid after the step counts are known (block.py:238-248)."""
	ys = _linspace2(up, ysteps)
	xs = _linspace2(right, xsteps)
	dg = (ys + p0)[:, None] + xs[None, :]
	dg = np.flip(dg, axis=-1)
	out = inverse_yx(dense, dg.reshape(-1, 2))
	return out.reshape(ysteps, xsteps, 2).astype(F32)


def line_steps(dense, p, right, up, xsteps=None, ysteps=None, xres=1, yres=1, column=None):
	"""Step counts of Line.warped_grid (block.py:226-236).

	The rough 2x2 grid is taken WITHOUT the column (the recursive call passes none) and is
	already fp32, so the extent is an fp32 difference; `xsteps * xres` with a Python float
	stays fp32 under NumPy >= 2 (xres rounded to fp32 first).
	"""
	_, _, _, xres_c = line_position(p, right, up, xres, column)
	if xsteps is None or ysteps is None:
		wx, wy = _rough_extent(dense, p, right, up)
	if xsteps is None:
		xsteps = max(2, int(np.ceil(wx * F32(xres_c))))
	if ysteps is None:
		ysteps = max(2, int(np.ceil(wy * F32(yres))))
	return xsteps, ysteps


def warped_grid(dense, p, right, up, xsteps=None, ysteps=None, xres=1, yres=1, column=None):
	"""Line.warped_grid (origami/core/block.py:223-248) -> (ysteps, xsteps, 2) fp32."""
	xsteps, ysteps = line_steps(dense, p, right, up, xsteps, ysteps, xres, yres, column)
	p0, right_c, up_c, _ = line_position(p, right, up, xres, column)
	return lattice_to_warped(dense, p0, right_c, up_c, xsteps, ysteps)


def dewarped_line_image(warped_page, dense, p, right, up, target_height=48, column=None):
	"""Line.dewarped_image (origami/core/block.py:250-259): remap + vertical flip."""
	g = warped_grid(dense, p, right, up, ysteps=target_height, column=column)
	return remap_linear(warped_page, g)[::-1, :]


def warped_path(dense, p, right, up, resolution=1):
	"""Line.warped_path (origami/core/block.py:261-267)."""
	g = warped_grid(dense, p, right, up, ysteps=2, xres=resolution)
	height = np.median(np.linalg.norm(g[1] - g[0], axis=-1))
	return np.mean(g, axis=0), abs(height)


def grid_bbox(g):
	"""A21: integer source-space box of an fp32 warped grid, Mask.__init__ convention
	(origami/core/mask.py:10-12): floor(min), ceil(max) as int32 -> (x0, y0, x1, y1)."""
	return (int(math.floor(g[..., 0].min())), int(math.floor(g[..., 1].min())),
		int(math.ceil(g[..., 0].max())), int(math.ceil(g[..., 1].max())))


# --------------------------------------------------------------------------------------
# A15  ConfidenceSampler.__call__     origami/batch/detect/lines.py:34-54
# --------------------------------------------------------------------------------------

def sample_label_counts(labels, page_shape, dense, p, right, up, n_classes, res=0.5):
	"""Half-resolution warped grid, scaled to the label map, NEAREST remap, bincount."""
	g = warped_grid(dense, p, right, up, xres=res, yres=res)
	h0, w0 = page_shape
	h1, w1 = labels.shape
	g = scale_map(g, w0, h0, w1, h1)
	lab = remap_nearest(labels, g)
	return np.bincount(lab.flatten(), minlength=n_classes)


# --------------------------------------------------------------------------------------
# A12  binarizers                      origami/core/binarize.py:11-43
# --------------------------------------------------------------------------------------

def _reflect_index(i, n):
	"""np.pad(mode="reflect") index map (no edge repeat), any overhang."""
	if n == 1:
		return np.zeros_like(i)
	period = 2 * (n - 1)
	i = np.mod(i, period)
	return np.where(i >= n, period - i, i)


def window_sums(pixels, w):
	"""Exact integer sums of p and p^2 over the centred w x w window, reflect-padded.

	skimage 0.16.2 _mean_std pads by (w//2+1, w//2) with mode="reflect", builds float64
	integral images and correlates with a (w+1)^2 corner kernel, which is the centred
	window sum; for 8-bit input every intermediate is an integer < 2^53, so it is exact.
	"""
	p = np.asarray(pixels).astype(np.int64)
	h, wd = p.shape
	r = w // 2
	yy = _reflect_index(np.arange(-r, h + r), h)
	xx = _reflect_index(np.arange(-r, wd + r), wd)
	pad = p[yy][:, xx]

	def box(a):
		c = np.cumsum(a, axis=0)
		c = np.vstack([np.zeros((1, a.shape[1]), np.int64), c])
		v = c[w:] - c[:-w]
		c = np.cumsum(v, axis=1)
		c = np.hstack([np.zeros((v.shape[0], 1), np.int64), c])
		return c[:, w:] - c[:, :-w]

	return box(pad), box(pad * pad)


def threshold_sauvola(pixels, window_size=15, k=0.2, r=None):
	"""skimage.filters.threshold_sauvola 0.16.2 for 8-bit input (float64).

	m = S1/w^2, g2 = S2/w^2, s = sqrt(clip(g2 - m*m, 0)), t = m*(1 + k*((s/r) - 1)),
	r = 127.5 for uint8 (0.5*(imax-imin) of the dtype limits).
	"""
	if window_size == 1 or window_size % 2 == 0:
		raise ValueError("Window size w = %s must be odd and greater than 1." % window_size)
	if r is None:
		r = 0.5 * (255 - 0)
	s1, s2 = window_sums(pixels, window_size)
	n = float(window_size ** 2)
	m = s1.astype(F64) / n
	g2 = s2.astype(F64) / n
	s = np.sqrt(np.clip(g2 - m * m, 0, None))
	return m * (1 + k * ((s / r) - 1))


def threshold_otsu(pixels):
	"""skimage.filters.threshold_otsu 0.16.2 for 8-bit input.

	Histogram over [min, max] with integer bin centres; float64 cumulative sums; first
	argmax of w1*w2*(m1-m2)^2.  A single-valued image raises ValueError, which
	origami/core/binarize.py:20-23 turns into the threshold 128.
	"""
	p = np.asarray(pixels)
	lo, hi = int(p.min()), int(p.max())
	if lo == hi:
		raise ValueError("single colour")
	hist = np.bincount(p.ravel(), minlength=hi + 1)[lo:].astype(F64)
	centers = np.arange(lo, hi + 1)
	w1 = np.cumsum(hist)
	w2 = np.cumsum(hist[::-1])[::-1]
	m1 = np.cumsum(hist * centers) / w1
	m2 = (np.cumsum((hist * centers)[::-1]) / w2[::-1])[::-1]
	var12 = w1[:-1] * w2[1:] * (m1[:-1] - m2[1:]) ** 2
	return int(centers[:-1][np.argmax(var12)])


def binarize(pixels, spec):
	"""binarize_with_threshold (origami/core/binarize.py:11-15): (pixels > t) * 255.

	spec: None | ("otsu",) | ("sauvola", window_size)
	"""
	pixels = np.asarray(pixels, dtype=np.uint8)
	if spec is None:
		return pixels
	if spec[0] == "otsu":
		try:
			t = threshold_otsu(pixels)
		except ValueError:
			t = 128
	elif spec[0] == "sauvola":
		t = threshold_sauvola(pixels, spec[1])
	else:
		raise ValueError(spec)
	return (pixels > t).astype(np.uint8) * 255


# --------------------------------------------------------------------------------------
# A18 / A19  label argmax and class masks   origami/core/predict.py:281-285, contours.py:488
# --------------------------------------------------------------------------------------

def argmax_labels(pr_masks):
	"""np.argmax(np.sum(pr_masks, axis=0), axis=-1).astype(uint8): fp32 sum over models in
	order, first maximum wins."""
	pr = np.asarray(pr_masks, dtype=F32)
	acc = pr[0].copy()
	for m in range(1, pr.shape[0]):
		acc = (acc + pr[m]).astype(F32)
	return np.argmax(acc, axis=-1).astype(np.uint8)


def class_masks(labels, values):
	"""pixels == c.value for every class value (origami/core/contours.py:488)."""
	return np.stack([(labels == v) for v in values]).astype(np.uint8)


# --------------------------------------------------------------------------------------
# A11  Line._extract_deskewed warpAffine stage   origami/core/block.py:155-172
# --------------------------------------------------------------------------------------

def affine_from_points(src_tri, dst_tri):
	"""cv2.getAffineTransform on fp32-rounded points, solved in float64 (LU in OpenCV;
	the tests compare the matrix with cv2's to 1e-9 and use cv2's matrix for pixel parity)."""
	src = np.asarray(src_tri, dtype=F32).astype(F64)
	dst = np.asarray(dst_tri, dtype=F32).astype(F64)
	A = np.zeros((6, 6))
	b = np.zeros(6)
	for i in range(3):
		A[i, 0:2] = src[i]
		A[i, 2] = 1
		A[i + 3, 3:5] = src[i]
		A[i + 3, 5] = 1
		b[i] = dst[i, 0]
		b[i + 3] = dst[i, 1]
	return np.linalg.solve(A, b).reshape(2, 3)


def warp_affine_linear(src, M, width, height, border=0):
	"""cv2.warpAffine(src u8, M, (width, height), INTER_LINEAR|INTER_AREA, BORDER_CONSTANT).

	imgwarp.cpp warpAffine: M inverted in double; per column adelta = rint(iM00*x*1024),
	bdelta = rint(iM10*x*1024); per row X0 = rint((iM01*y + iM02)*1024) + 16 (half of
	1024/32), likewise Y0; X = (X0 + adelta) >> 5 gives ix = X >> 5, fx = X & 31; then the
	same integer bilinear blend as remap.
	"""
	src = np.asarray(src, dtype=np.uint8)
	M = np.asarray(M, dtype=F64)
	D = M[0, 0] * M[1, 1] - M[0, 1] * M[1, 0]
	D = 1.0 / D if D != 0 else 0.0
	A11 = M[1, 1] * D
	A22 = M[0, 0] * D
	i00, i01, i10, i11 = A11, M[0, 1] * (-D), M[1, 0] * (-D), A22
	b1 = -i00 * M[0, 2] - i01 * M[1, 2]
	b2 = -i10 * M[0, 2] - i11 * M[1, 2]
	xs = np.arange(width, dtype=F64)
	ys = np.arange(height, dtype=F64)
	adelta = np.rint(i00 * xs * 1024).astype(np.int64)
	bdelta = np.rint(i10 * xs * 1024).astype(np.int64)
	X0 = np.rint((i01 * ys + b1) * 1024).astype(np.int64) + 16
	Y0 = np.rint((i11 * ys + b2) * 1024).astype(np.int64) + 16
	X = (X0[:, None] + adelta[None, :]) >> 5
	Y = (Y0[:, None] + bdelta[None, :]) >> 5
	ix = _sat_short(X >> 5)
	iy = _sat_short(Y >> 5)
	fx = X & 31
	fy = Y & 31
	p00 = _tap(src, iy, ix, border)
	p01 = _tap(src, iy, ix + 1, border)
	p10 = _tap(src, iy + 1, ix, border)
	p11 = _tap(src, iy + 1, ix + 1, border)
	acc = ((32 - fx) * (32 - fy) * p00 + fx * (32 - fy) * p01
		+ (32 - fx) * fy * p10 + fx * fy * p11)
	return ((acc + 512) >> 10).astype(np.uint8)


# --------------------------------------------------------------------------------------
# A20  Regions.binarized            origami/batch/detect/layout.py:157-186
# --------------------------------------------------------------------------------------

def dilate3(mask):
	"""skimage.morphology.binary_dilation(mask, square(3)): 3x3 maximum, False outside."""
	m = np.asarray(mask, dtype=bool)
	p = np.pad(m, 1, mode="constant", constant_values=False)
	out = np.zeros_like(m)
	for dy in range(3):
		for dx in range(3):
			out |= p[dy:dy + m.shape[0], dx:dx + m.shape[1]]
	return out


def regions_binarized(page, grid, s, median_line_height, separator_labels=()):
	"""Regions.binarized: Sauvola (window from the median line height) on the dewarped page, OR the
	dilated NEAREST-dewarped separator masks, as float32."""
	gray = dewarp_page(page, grid, s)
	window_size = int(median_line_height) // 2
	if window_size % 2 == 0:
		window_size += 1
	window_size = max(window_size, 3)
	binary = gray > threshold_sauvola(gray, window_size)
	for labels, bg in separator_labels:
		mask = (np.asarray(labels) != bg).astype(np.uint8) * 255
		mask = dewarp_image(mask, (page.shape[1], page.shape[0]), grid, s, INTER_NEAREST)
		binary = np.logical_or(binary, dilate3(mask > 0))
	return binary.astype(np.float32)


# --------------------------------------------------------------------------------------
# (f)4  OCR hand-off                origami/batch/detect/ocr.py:238-268
# --------------------------------------------------------------------------------------

def ocr_batch(strips, height, w_max, min_width=6, min_height=6, pad_value=0):
	"""The strips the OCR stage keeps (`im.width >= min_width and im.height >= min_height`,
	ocr.py:241-246), in order, as the padded batch its predictor's input pipeline forms: every kept
	image in the top-left corner of a (height, w_max) slot filled with pad_value.
	Returns (batch (n_kept, height, w_max) u8, widths, kept indices)."""
	kept = [i for i, s in enumerate(strips) if s.shape[1] >= min_width and s.shape[0] >= min_height
		and s.shape[0] <= height and s.shape[1] <= w_max]
	batch = np.full((len(kept), height, w_max), pad_value, dtype=np.uint8)
	for j, i in enumerate(kept):
		s = strips[i]
		batch[j, :s.shape[0], :s.shape[1]] = s
	return batch, np.array([strips[i].shape[1] for i in kept], dtype=np.int32), np.array(kept, dtype=np.int32)


# --------------------------------------------------------------------------------------
# (f)2  Grid.transformer            origami/core/dewarp.py:143-161, core/lingrid.py:192-301
# --------------------------------------------------------------------------------------

def _ray_box_hit(p, direction, box, size):
	"""First crossing of the ray p + t * direction * size (t in [0, 1]) with the outline of
	box = (minx, miny, maxx, maxy) -- what `ray.intersection(LinearRing(box))` gives in
	lingrid.py:62-78 for a point inside the box (shapely absent: restated, slab method)."""
	d = np.asarray(direction, F64) * size
	cands = []
	for axis, lo, hi in ((0, box[0], box[2]), (1, box[1], box[3])):
		for bound in (lo, hi):
			if d[axis] != 0:
				t = (bound - p[axis]) / d[axis]
				if 0 <= t <= 1:
					q = np.array(p, F64) + t * d
					q[axis] = bound
					o = 1 - axis
					if box[o] <= q[o] <= box[o + 2]:
						cands.append((t, float(q[0]), float(q[1])))
	if not cands:
		raise ValueError("point outside given domain")
	return min(cands)[1:]


def lininterp_points(points, values, bounds, margin=1):
	"""The extrapolation point set of lingrid.InterpolatorFactory (lingrid.py:207-262): hull
	vertices, their projections on the enlarged box along the hull-edge normals, the corners."""
	import scipy.spatial
	points = np.asarray(points, F64)
	values = np.asarray(values, F64).reshape(len(points), -1)
	minx, miny, maxx, maxy = bounds
	box = (minx - margin, miny - margin, maxx + margin, maxy + margin)
	size = max(maxx - minx, maxy - miny) + 2 * margin + 1
	hull = scipy.spatial.ConvexHull(points)
	hv = list(hull.vertices)
	pts = [points[i] for i in hv]
	vals = [values[i] for i in hv]
	on = {"L": [], "R": [], "T": [], "B": []}
	proj = []
	for a_i, b_i in zip(hv, hv[1:] + hv[:1]):
		a, b = points[a_i], points[b_i]
		v = a - b
		if np.linalg.norm(v) < 1e-2:
			continue
		nrm = np.array([-v[1], v[0]]) / np.linalg.norm(v)
		for q, val in ((a, values[a_i]), (b, values[b_i])):
			x, y = _ray_box_hit(q, nrm, box, size)
			e = (x, y, val)
			proj.append(e)
			if x == box[0]:
				on["L"].append(e)
			elif x == box[2]:
				on["R"].append(e)
			if y == box[1]:
				on["T"].append(e)
			elif y == box[3]:
				on["B"].append(e)
	for x, y, val in proj:
		pts.append(np.array([x, y]))
		vals.append(val)

	def corner(cx, cy, p1, p2):
		if p1 is None and p2 is None:
			return
		if p1 is None:
			val = p2[2]
		elif p2 is None:
			val = p1[2]
		else:
			a, b = np.array(p1[:2]), np.array(p2[:2])
			dt = np.linalg.norm(b - a)
			if dt == 0:
				return
			# foot of the perpendicular from the corner onto the line a b (sympy Line.projection, lingrid.py:115-120)
			t = ((cx - a[0]) * (b[0] - a[0]) + (cy - a[1]) * (b[1] - a[1])) / (dt * dt)
			foot = a + t * (b - a)
			w = np.linalg.norm(foot - a) / dt
			val = (1 - w) * p1[2] + w * p2[2]
		pts.append(np.array([cx, cy]))
		vals.append(val)

	pick = lambda lst, f, k: f(lst, key=lambda e: e[k]) if lst else None  # noqa: E731
	corner(box[0], box[1], pick(on["L"], min, 1), pick(on["T"], min, 0))
	corner(box[2], box[1], pick(on["R"], min, 1), pick(on["T"], max, 0))
	corner(box[2], box[3], pick(on["R"], max, 1), pick(on["B"], max, 0))
	corner(box[0], box[3], pick(on["L"], max, 1), pick(on["B"], min, 0))
	return np.array(pts), np.array(vals)


def lininterp(points, values, bounds):
	"""lingrid.lininterp (lingrid.py:173-189, 266-275, 300-301) evaluated with SciPy's own
	LinearNDInterpolator, as the reference does: clip to the bounds, interpolate inside the hull of
	the data, use the extrapolating interpolator where that gives NaN."""
	import scipy.interpolate
	points = np.asarray(points, F64)
	values = np.asarray(values, F64).reshape(len(points), -1)
	inter = scipy.interpolate.LinearNDInterpolator(points, values, fill_value=np.nan)
	ep, ev = lininterp_points(points, values, bounds)
	extra = scipy.interpolate.LinearNDInterpolator(ep, ev, fill_value=np.nan)

	def f(pts):
		pts = np.array(pts, dtype=F64).reshape(-1, 2)
		pts[:, 0] = np.clip(pts[:, 0], bounds[0], bounds[2])
		pts[:, 1] = np.clip(pts[:, 1], bounds[1], bounds[3])
		ri, rx = inter(pts), extra(pts)
		return np.where(np.isnan(ri), rx, ri), np.isnan(ri).any(axis=1)

	return f


def transformer(dense, s):
	"""Grid.transformer (core/dewarp.py:555-559 + Transformer :143-161) -> f(pts (n, 2)) ->
	(dewarped (n, 2), outside-the-hull flags)."""
	nodes = np.asarray(dense)[::s, ::s]
	h, w = nodes.shape[:2]
	source = nodes.reshape(h * w, 2).astype(F64)
	target = np.flip(np.dstack(np.mgrid[0:h, 0:w]), axis=-1).reshape(h * w, 2) * s
	b = (source[:, 0].min(), source[:, 1].min(), source[:, 0].max(), source[:, 1].max())
	return lininterp(source, target.astype(F64), b)


# --------------------------------------------------------------------------------------
# A11  polygon masks                 origami/core/mask.py:8-74
# --------------------------------------------------------------------------------------

def _clip_ring(ring, axis, bound, keep_greater):
	"""Sutherland-Hodgman: the part of a closed ring (list of points) on one side of an axis-parallel line."""
	out = []
	n = len(ring)
	for i in range(n):
		a, b = ring[i], ring[(i + 1) % n]
		ina = a[axis] >= bound if keep_greater else a[axis] <= bound
		inb = b[axis] >= bound if keep_greater else b[axis] <= bound
		if ina:
			out.append(a)
		if ina != inb:
			t = (bound - a[axis]) / (b[axis] - a[axis])
			out.append((a[0] + t * (b[0] - a[0]), a[1] + t * (b[1] - a[1])))
	return out


def polygon_coverage(ring, w, h):
	"""Exact area of polygon /\\ pixel square for every pixel of a (h, w) raster, by clipping the
	ring against the four sides of each pixel and the shoelace formula (pure Python: small cases).
	ring: (k, 2) vertices, closing edge implied."""
	pts = [(float(x), float(y)) for x, y in np.asarray(ring, F64).reshape(-1, 2)]
	if len(pts) > 1 and pts[0] == pts[-1]:
		pts = pts[:-1]
	cov = np.zeros((h, w), F64)
	xs = [p[0] for p in pts]
	ys = [p[1] for p in pts]
	for y in range(max(0, int(math.floor(min(ys)))), min(h, int(math.ceil(max(ys))))):
		row = _clip_ring(_clip_ring(pts, 1, y, True), 1, y + 1, False)
		if len(row) < 3:
			continue
		for x in range(max(0, int(math.floor(min(p[0] for p in row)))), min(w, int(math.ceil(max(p[0] for p in row))))):
			cell = _clip_ring(_clip_ring(row, 0, x, True), 0, x + 1, False)
			if len(cell) >= 3:
				a = 0.0
				for i in range(len(cell)):
					p, q = cell[i], cell[(i + 1) % len(cell)]
					a += p[0] * q[1] - q[0] * p[1]
				cov[y, x] = abs(a) / 2
	return cov


def polygon_mask(ring, w, h):
	"""Mask.__init__ (mask.py:36-43) for one polygon whose vertices are already relative to the mask
	origin: vertices rounded to integers, a pixel is set where cairo's 8-bit coverage is non-zero
	(coverage >= 1/510).  cairo's rasteriser is not available: exact coverage stands in (unpinned)."""
	pts = np.round(np.asarray(ring, F64).reshape(-1, 2)).astype(np.int32)
	return polygon_coverage(pts, w, h) >= 0.5 / 255.0


# --------------------------------------------------------------------------------------
# (f)3  SplitDetector               origami/batch/detect/layout.py:915-944
# --------------------------------------------------------------------------------------

def split_detector(pixels, scale, quantile=0.9, smooth=1, intensity=0.05, width=2, border=0.1):
	"""SplitDetector.__call__ with the reference's own SciPy / NumPy calls (they are installed here).
	Returns (peaks, info, column quantiles before smoothing)."""
	import scipy.ndimage
	import scipy.signal
	if pixels.dtype == np.uint8:
		pixels = pixels.astype(np.float32) / 255.0
	assert pixels.dtype == np.float32
	freq, dens = scipy.signal.periodogram(pixels, axis=0)
	colq = np.quantile(dens, quantile, axis=0)
	ink_h = scipy.ndimage.convolve(colq, np.ones(int(smooth * scale)) / int(smooth * scale), mode="nearest")
	span = int(border * len(ink_h))
	ink_h[:span] = 0
	ink_h[-span:] = 0
	peaks, info = scipy.signal.find_peaks(-ink_h, height=-intensity, distance=int(width * scale))
	return peaks, info, colq


# --------------------------------------------------------------------------------------
# (f)1  GridFactory                 origami/core/dewarp.py:108-141, 176-263, 292-493
# --------------------------------------------------------------------------------------

def _field(points, values, size, phi0):
	"""Field (dewarp.py:108-122): angles by lininterp, directions as float32 cos / sin."""
	if len(points) <= 2:
		angle = lambda pts: np.full(len(pts), phi0)  # noqa: E731
	else:
		f = lininterp(np.asarray(points, F64), np.asarray(values, F64), (0, 0, size[0], size[1]))
		angle = lambda pts: f(np.array(pts))[0][:, 0]  # noqa: E731

	def get(pts):
		a = angle(pts)
		return np.squeeze(np.dstack([np.cos(a, dtype=F32), np.sin(a, dtype=F32)]), axis=0)
	return get


def _estimate_extent(get, size, axis, limit, step_size):
	"""Field.estimate_extent (dewarp.py:123-141)."""
	pts = np.array([[0, y] for y in range(0, size[1 - axis], step_size)]).astype(F32)
	if axis != 0:
		pts = np.flip(pts, axis=-1).copy()
	n_steps = 1
	max_steps = 2 * (1 + size[axis] // n_steps)
	while np.any(pts[:, axis] < limit) and n_steps < max_steps:
		pts += get(pts) * step_size
		n_steps += 1
	if n_steps >= max_steps:
		raise RuntimeError("n_steps exceeded")
	return n_steps


def _nearest_crossing(p0, p1, line):
	"""Nearest point to p0 where segment p0-p1 crosses the polyline `line` (k, 2), or None:
	ShapelyBatchIntersections.__call__ (dewarp.py:218-262) for one ray and one row."""
	a = np.asarray(p0, F64)
	r = np.asarray(p1, F64) - a
	q = line[:-1].astype(F64)
	s = line[1:].astype(F64) - q
	den = r[0] * s[:, 1] - r[1] * s[:, 0]
	with np.errstate(divide="ignore", invalid="ignore"):
		t = ((q[:, 0] - a[0]) * s[:, 1] - (q[:, 1] - a[1]) * s[:, 0]) / den
		u = ((q[:, 0] - a[0]) * r[1] - (q[:, 1] - a[1]) * r[0]) / den
	ok = (den != 0) & (t >= 0) & (t <= 1) & (u >= 0) & (u <= 1)
	if not ok.any():
		return None
	pts = a[None, :] + t[ok, None] * r[None, :]
	return pts[np.argmin(np.linalg.norm(pts - a, axis=1))]


def grid_factory(size, samples_h, samples_v, grid_res=25, num_threads=2):
	"""GridFactory.grid_hv (dewarp.py:292-493) for a page of `size` = (width, height); samples_* =
	(points (n, 2), angles (n,)).  Returns the float32 (gh, gw, 2) grid."""
	width, height = size
	get_h = _field(samples_h[0], samples_h[1], size, 0.0)
	get_v = _field(samples_v[0], samples_v[1], size, math.pi / 2)
	gw = _estimate_extent(get_h, size, 0, width, grid_res)
	gh = _estimate_extent(get_v, size, 1, height, grid_res)
	grid_h = np.zeros((gh, gw, 2), dtype=F32)
	pts = np.array([[0, gy * grid_res] for gy in range(gh)], dtype=F32)
	for gx in range(gw):
		grid_h[:, gx, :] = pts
		pts += get_h(pts) * grid_res
	# make_row_slice (dewarp.py:198-216) pushes the end points of rows 1.. outwards, in place
	large = gw * grid_res
	for gy in range(1, gh):
		for a, b in ((1, 0), (-2, -1)):
			v = grid_h[gy, b] - grid_h[gy, a]
			v /= np.linalg.norm(v)
			grid_h[gy, b] = grid_h[gy, b] + large * v
	grid_hv = np.zeros(grid_h.shape, dtype=F32)
	s = math.ceil(gw / num_threads) if num_threads >= 2 else gw
	for c0 in range(0, gw, s):
		sel = slice(c0, c0 + s)
		max_dy = 0
		for gy in range(1, gh):
			max_dy = max(max_dy, np.max(grid_h[gy, sel, 1]) - np.min(grid_h[gy - 1, sel, 1]))
		max_step_len = max_dy / math.cos(60 * (math.pi / 180))
		pts0 = grid_h[0][sel].copy()
		for gy in range(gh - 1):
			grid_hv[gy, sel, :] = pts0
			pts1 = pts0 + get_v(pts0) * max_step_len
			for i in range(len(pts0)):
				hit, dy = None, 0
				while hit is None:
					if gy + 1 + dy >= gh:
						raise RuntimeError("failed to find intersection")
					hit = _nearest_crossing(pts0[i], pts1[i], grid_h[gy + 1 + dy])
					dy += 1
				pts1[i] = hit
			pts0 = pts1
		grid_hv[-1, sel, :] = pts0

	def extend(grid, get, start, cond, step, stack, first):
		borders = [start]
		while cond(borders[-1]):
			borders.append((borders[-1] + get(borders[-1]) * step).astype(F32))
		new = borders[1:]
		if not new:
			return grid
		return stack(new, first)

	gh0 = grid_hv.shape[0]
	new = extend(grid_hv, get_h, grid_hv[:, 0].copy(), lambda x: np.any(x[:, 0] > 0), -grid_res,
		lambda n, g: np.hstack([b.reshape(gh0, 1, 2) for b in reversed(n)] + [g]), grid_hv)
	grid_hv = new
	grid_hv = extend(grid_hv, get_h, grid_hv[:, -1].copy(), lambda x: np.any(x[:, 0] < width), grid_res,
		lambda n, g: np.hstack([g] + [b.reshape(gh0, 1, 2) for b in n]), grid_hv)
	gw1 = grid_hv.shape[1]
	grid_hv = extend(grid_hv, get_v, grid_hv[0, :].copy(), lambda x: np.any(x[:, 1] > 0), -grid_res,
		lambda n, g: np.vstack([b.reshape(1, gw1, 2) for b in reversed(n)] + [g]), grid_hv)
	grid_hv = extend(grid_hv, get_v, grid_hv[-1, :].copy(), lambda x: np.any(x[:, 1] < height), grid_res,
		lambda n, g: np.vstack([g] + [b.reshape(1, gw1, 2) for b in n]), grid_hv)
	return grid_hv
