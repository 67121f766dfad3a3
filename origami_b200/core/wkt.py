"""
The two WKT forms Origami's artifacts hold (contours.zip, lines.zip: POLYGON with optional holes,
LINESTRING; reference docs/formats.md) as plain coordinate rings.  The reference parses them with
shapely (shapely.wkt.loads); shapely is not installed here and only the coordinates are needed.
"""

import re

import numpy as np

_RING = re.compile(r"\(([^()]*)\)")


def parse_wkt(text):
	"""'POLYGON ((x y, ...), (hole ...))' | 'LINESTRING (x y, ...)' -> (kind, [ (k, 2) float64 rings ])."""
	text = text.strip()
	kind = text.split("(", 1)[0].strip().upper()
	if kind not in ("POLYGON", "LINESTRING"):
		raise ValueError("unsupported geometry %r" % kind)
	rings = []
	for m in _RING.finditer(text):
		body = m.group(1).strip()
		if not body:
			continue
		rings.append(np.array([[float(v) for v in pt.split()[:2]] for pt in body.split(",")], dtype=np.float64))
	if not rings:
		raise ValueError("empty geometry")
	return kind, rings


def format_wkt(kind, rings):
	def ring(r):
		return "(" + ", ".join("%s %s" % (repr(float(x)), repr(float(y))) for x, y in r) + ")"
	if kind == "LINESTRING":
		return "LINESTRING " + ring(rings[0])
	return "POLYGON (" + ", ".join(ring(r) for r in rings) + ")"


def ring_area(r):
	"""Signed shoelace area of a closed ring."""
	x, y = r[:, 0], r[:, 1]
	return 0.5 * float(np.dot(x[:-1], y[1:]) - np.dot(x[1:], y[:-1])) if len(r) > 2 else 0.0


def polygon_area(rings):
	"""shapely's Polygon.area: exterior minus holes."""
	return abs(ring_area(rings[0])) - sum(abs(ring_area(h)) for h in rings[1:])


