"""Geometry helper with the interface of the reference's origami/core/math.py:69-91."""

import numpy as np


class Geometry:
	def __init__(self, w, h):
		self._size = (w, h)

	@property
	def size(self):
		return self._size

	@property
	def area(self):
		w, h = self.size
		return w * h

	@property
	def diameter(self):
		w, h = self.size
		return np.sqrt(w * w + h * h)

	def rel_length(self, x):
		return self.diameter * x

	def rel_area(self, a):
		return (self.diameter * a) ** 2
