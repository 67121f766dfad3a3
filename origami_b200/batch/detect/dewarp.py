"""
Contour dewarping of the dewarp stage (reference origami/batch/detect/dewarp.py:16-39): every
vertex of every contour of contours.zip goes through Grid.transformer.  The reference calls
shapely.ops.transform per geometry, which evaluates the interpolator once per ring; here the
vertices of ALL geometries of a page are gathered and transformed in one kernel launch.

shapely is not installed, so the geometries are handled as plain coordinate rings: the two WKT
forms contours.zip holds (POLYGON with optional holes, LINESTRING; docs/formats.md) are parsed and
written here, areas come from the shoelace formula.  What needs a geometry engine -- repairing an
invalid polygon with buffer(0) -- is not done: such a polygon is reported and passed through.
"""

import logging

import numpy as np

from ...core.wkt import format_wkt, parse_wkt, polygon_area, ring_area  # noqa: F401


def dewarp_geometries(transformer, geometries):
	"""[(kind, rings)] -> the same structure with every vertex sent through the transformer
	(one launch for all vertices of all geometries)."""
	sizes = [len(r) for _, rings in geometries for r in rings]
	if not sizes:
		return []
	allpts = np.concatenate([r for _, rings in geometries for r in rings], axis=0)
	out = transformer.transform_points(allpts)
	res, k = [], 0
	for kind, rings in geometries:
		new = []
		for r in rings:
			new.append(out[k:k + len(r)])
			k += len(r)
		res.append((kind, new))
	return res


def dewarped_contours(named_wkt, transformer, min_areas):
	"""reference dewarp.py:16-39 on (name, wkt bytes/str) pairs as read from contours.zip; yields
	(name, wkt bytes) of the dewarped contours, dropping those that shrink below the minimum area
	of their predictor (min_areas keyed by the first path component)."""
	items = []
	for name, data in named_wkt:
		if not name.endswith(".wkt"):
			continue
		text = data.decode("utf8") if isinstance(data, (bytes, bytearray)) else data
		items.append((name, parse_wkt(text)))
	moved = dewarp_geometries(transformer, [g for _, g in items])
	for (name, (kind, rings)), (_, new) in zip(items, moved):
		path = tuple(name.rsplit(".", 1)[0].split("/"))
		if kind == "POLYGON":
			area = polygon_area(new)
			if area < min_areas.get(path[0], 0):
				logging.warning("lost contour %s (A=%.1f) during dewarping." % (path, polygon_area(rings)))
				continue
		yield name, format_wkt(kind, new).encode("utf8")
