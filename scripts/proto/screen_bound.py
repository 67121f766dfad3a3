"""Numerical check of the error bound used by the fp32 screening of the line kernel:
delta = fma(a, fma(b, d11 - d01, d10), b * d01) in float32 against the exact bilinear offset."""
import numpy as np
from fractions import Fraction

rng = np.random.default_rng(1)
f32 = np.float32

def fma32(a, b, c):
	return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)

worst = 0.0
for scale, base in [(1, 4000), (1, 0.5), (4, 20000), (0.01, 100), (100, 5000), (1, 30000)]:
	n = 2_000_000
	v00 = (base + rng.standard_normal(n) * scale).astype(f32)
	v01 = (v00 + rng.standard_normal(n) * scale).astype(f32)
	v10 = (v00 + rng.standard_normal(n) * scale).astype(f32)
	v11 = (v10 + rng.standard_normal(n) * scale).astype(f32)
	A = rng.random(n)
	B = rng.random(n)
	a = (np.floor(A * 2**23) / 2**23).astype(f32)
	b = (np.floor(B * 2**23) / 2**23).astype(f32)
	d01 = v01 - v00
	d10 = v10 - v00
	d11 = v11 - v10
	dd = d11 - d01
	inner = fma32(b, dd, d10)
	m = b * d01
	delta = fma32(a, inner, m)
	G = np.maximum(np.maximum(np.abs(d01), np.abs(d10)), np.abs(d11))
	# exact (float64 is plenty: the inputs are float32, errors ~1e-16 relative)
	V00, V01, V10, V11 = (x.astype(np.float64) for x in (v00, v01, v10, v11))
	ref = (V01 - V00) * B * (1 - A) + (V10 - V00) * A * (1 - B) + (V11 - V00) * A * B
	err = np.abs(delta.astype(np.float64) - ref)
	ratio = err / np.maximum(G.astype(np.float64) * 2.0**-24, 1e-300)
	worst = max(worst, ratio.max())
	print(scale, base, "max err / (u G) = %.2f" % ratio.max(), " mean %.2f" % ratio.mean())
print("worst", worst, "(bound used: 40)")
