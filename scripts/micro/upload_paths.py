"""How fast can a pageable 4800x7200 uint8 NumPy page reach the device? (development aid)
current chunked staging (ops.to_device_image) vs a plain pageable copy vs cudaHostRegister in place."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from origami_b200 import ops, synth

dev = torch.device("cuda:0")
page = synth.make_page(4800, 7200, seed=0, kind="text")
pages = [page.copy() for _ in range(4)]
rt = torch.cuda.cudart()

def t(fn, n=8):
	fn(pages[0]); torch.cuda.synchronize()
	t0 = time.perf_counter()
	for i in range(n):
		fn(pages[i % 4])
	torch.cuda.synchronize()
	return (time.perf_counter() - t0) / n * 1e3

dst = ops.alloc_image(7200, 4800, dev)
print("ops.to_device_image          %.2f ms" % t(lambda a: ops.to_device_image(a, dev)))
print("torch.from_numpy(a).to(dev)  %.2f ms" % t(lambda a: torch.from_numpy(a).to(dev)))
def reg(a):
	ta = torch.from_numpy(a)
	rc = rt.cudaHostRegister(ta.data_ptr(), ta.numel(), 0)
	dst.copy_(ta, non_blocking=True)
	torch.cuda.synchronize()
	rt.cudaHostUnregister(ta.data_ptr())
print("register + copy + unregister %.2f ms" % t(reg))
def regonly(a):
	ta = torch.from_numpy(a)
	rt.cudaHostRegister(ta.data_ptr(), ta.numel(), 0)
	rt.cudaHostUnregister(ta.data_ptr())
print("register + unregister only   %.2f ms" % t(regonly))
pin = torch.empty((7200, 4800), dtype=torch.uint8).pin_memory()
def staged(a):
	pin.numpy()[...] = a
	dst.copy_(pin, non_blocking=True)
print("one memcpy to pinned + copy  %.2f ms" % t(staged))
def h2d_only(a):
	dst.copy_(pin, non_blocking=True)
print("pinned copy alone            %.2f ms" % t(h2d_only))
import concurrent.futures
src = pages[0].reshape(-1)
dstn = pin.numpy().reshape(-1)
for nt in (1, 2, 3, 4, 6, 8, 16):
	ops._COPY_THREADS = nt
	ops._COPY_POOL = concurrent.futures.ThreadPoolExecutor(max_workers=nt) if nt > 1 else None
	def pc(a):
		if nt == 1:
			np.copyto(dstn, a.reshape(-1))
		else:
			ops.parallel_copy(dstn, a.reshape(-1))
	print("host copy 35 MB, %2d threads  %.2f ms" % (nt, t(pc)))
	print("   to_device_image          %.2f ms" % t(lambda a: ops.to_device_image(a, dev)))
for nt in (3, 4, 5):
	ops._COPY_THREADS = nt
	ops._COPY_POOL = concurrent.futures.ThreadPoolExecutor(max_workers=nt)
	for ch in (1, 2, 3, 4, 6, 8):
		ops._STAGE_CHUNKS = ch
		print("threads %d chunks %d  to_device_image %.2f ms" % (nt, ch, t(lambda a: ops.to_device_image(a, dev), n=12)))
