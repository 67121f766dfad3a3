"""
Builds liborigami_b200.so (the C-ABI library declared in include/origami_b200.h) with nvcc
for sm_100a, in-tree, so the built library travels with the repository snapshot.

    python -m origami_b200.build [--force]
"""

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "liborigami_b200.so")

SOURCES = ["og_api.cu", "og_page.cu", "og_lines.cu", "og_strips.cu", "og_binarize.cu", "og_sauvola_warp.cu", "og_segment.cu", "og_handoff.cu", "og_geometry.cu"]

NVCC_FLAGS = [
	"-gencode", "arch=compute_100a,code=sm_100a",
	"-O3", "-lineinfo", "-std=c++17",
	"--fmad=false",  # parity: no implicit FMA contraction anywhere (explicit *_rn are used on top)
	"-Xcompiler", "-fPIC", "-Xcompiler", "-O2",
	"-shared",
]


def _nvcc():
	for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
		if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
			return cand
	return "nvcc"


def needs_build():
	if not os.path.exists(LIB_PATH):
		return True
	t = os.path.getmtime(LIB_PATH)
	deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
	deps.append(os.path.join(HERE, "..", "include", "origami_b200.h"))
	return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
	if not force and not needs_build():
		return LIB_PATH
	os.makedirs(LIB_DIR, exist_ok=True)
	cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
	cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB_PATH]
	res = subprocess.run(cmd, capture_output=True, text=True)
	if res.returncode != 0:
		raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
	if verbose:
		print(res.stderr)
	return LIB_PATH


if __name__ == "__main__":
	print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
