"""
Segmentation artifacts with the interface of the reference's origami/core/segment.py:13-86 and
origami/core/predict.py:23-40, 123-146, 371-401: segment.zip holds, per predictor, the label map as
a paletted PNG plus a JSON with the class table (docs/formats.md:40-70).  This is the wire format on
the input side of the segmentation post-process (label maps for ConfidenceSampler, separator masks
for Regions.binarized) and on its output side (labels produced by og_argmax_labels / og_argmax_tile).
Host-side codec work: PIL decodes / encodes the PNG, the label map then lives on the GPU.
"""

import enum
import io
import json
import zipfile

import numpy as np
import PIL.Image

from .. import ops

# matplotlib's "tab10" palette (the reference reads it from matplotlib, predict.py:20-22)
_TAB10 = np.array([
	(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189),
	(140, 86, 75), (227, 119, 194), (127, 127, 127), (188, 189, 34), (23, 190, 207)], dtype=np.float64)


class RegionLabel(enum.Enum):
	pass


class SeparatorLabel(enum.Enum):
	pass


class PredictorType(enum.Enum):
	REGION = 1
	SEPARATOR = 2

	def classes(self, c):
		if self == PredictorType.REGION:
			return enum.Enum("RegionLabel", c, type=RegionLabel)
		elif self == PredictorType.SEPARATOR:
			return enum.Enum("SeparatorLabel", c, type=SeparatorLabel)
		else:
			raise ValueError(self)


def colorize(labels, background=None):
	"""predict.py:25-40: the label map as a paletted image, background label painted white."""
	colors = _TAB10.flatten().copy()
	if background:
		colors[background * 3:background * 3 + 3] = 255
	im = PIL.Image.fromarray(np.ascontiguousarray(labels, dtype=np.uint8), "P")
	pil_pal = np.zeros((768,), dtype=np.uint8)
	pil_pal[:len(colors)] = colors
	im.putpalette(pil_pal)
	return im


class Prediction:
	def __init__(self, type, name, labels, classes):
		self._type = type
		self._name = name
		self._labels = labels
		self._classes = classes
		self._background = self._classes["BACKGROUND"]

	@property
	def type(self):
		return self._type

	@property
	def name(self):
		return self._name

	@property
	def background_label(self):
		return self._background

	@property
	def labels(self):
		return self._labels

	@property
	def classes(self):
		return self._classes

	@property
	def colorized(self):
		return colorize(self._labels, self.background_label.value)

	def device_labels(self, device=None):
		"""The label map as a pitched CUDA image (input of og_sample_labels / og_label_mask_ne)."""
		return ops.to_device_image(np.ascontiguousarray(self._labels, dtype=np.uint8), device)


class Segmentation:
	def __init__(self, predictions):
		self._predictions = tuple(predictions)

	@property
	def predictions(self):
		return self._predictions

	@property
	def size(self):
		shape = self._predictions[0].labels.shape
		return tuple(reversed(list(shape)[:2]))

	@staticmethod
	def open(path):
		predictions = []
		with zipfile.ZipFile(path, "r") as zf:
			tasks = [name.rsplit(".", 1)[0] for name in zf.namelist() if name.endswith(".png")]
			for task in tasks:
				with io.BytesIO(zf.read(task + ".png")) as f:
					im = PIL.Image.open(f)
					im.load()
				meta = json.loads(zf.read(task + ".json"))
				t = PredictorType[meta["type"]]
				classes = t.classes(meta["classes"])
				predictions.append(Prediction(t, meta["name"], np.array(im), classes))
		return Segmentation(predictions)

	def save(self, path):
		with zipfile.ZipFile(path, "w") as zf:
			for p in self._predictions:
				with io.BytesIO() as f:
					p.colorized.save(f, "png", optimize=True)
					zf.writestr("%s.png" % p.name, f.getvalue())
				meta = dict(
					type=p.type.name,
					name=p.name,
					classes=dict([(m.name, m.value) for m in p.classes]))
				zf.writestr("%s.json" % p.name, json.dumps(meta))


def write_line_images(zf, images, flat=False):
	"""origami/batch/export/lines.py:57-67: every (path, image) of LineExtractor(...)(lines) as an
	optimised PNG inside an open zipfile.  `image` may be a PIL image or a uint8 ndarray."""
	zip_sep = "-" if flat else "/"
	for stem, im in images:
		if not isinstance(im, PIL.Image.Image):
			im = PIL.Image.fromarray(im)
		with io.BytesIO() as f:
			im.save(f, format="png", optimize=True)
			data = f.getvalue()
		zf.writestr("%s.png" % zip_sep.join(stem), data)
