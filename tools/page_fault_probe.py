"""How expensive is the first touch of freshly allocated host memory on this box, and do transparent huge pages or
MADV_POPULATE_WRITE help?  (The D2H unpack of a product into a caller-allocated matrix pays exactly this.)"""
import ctypes
import mmap
import time

import numpy as np

libc = ctypes.CDLL("libc.so.6", use_errno=True)
libc.madvise.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
MADV_HUGEPAGE, MADV_POPULATE_WRITE = 14, 23
try:
    print("thp:", open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip(),
          "| defrag:", open("/sys/kernel/mm/transparent_hugepage/defrag").read().strip())
except OSError as e:
    print("thp: unreadable", e)
n = 4096 * 4096
src = np.random.default_rng(0).standard_normal(n)
for mode in ("plain", "hugepage", "populate", "hugepage+populate", "plain"):
    dst = np.empty(n)   # fresh mmap, untouched
    addr, nbytes = dst.ctypes.data, dst.nbytes
    lo = (addr + (2 << 20) - 1) & ~((2 << 20) - 1)
    t0 = time.perf_counter()
    if "hugepage" in mode:
        rc = libc.madvise(lo, nbytes - (lo - addr), MADV_HUGEPAGE)
        if rc != 0:
            print(mode, "madvise(HUGEPAGE) errno", ctypes.get_errno())
    if "populate" in mode:
        page = addr & ~4095
        rc = libc.madvise(page, nbytes + (addr - page), MADV_POPULATE_WRITE)
        if rc != 0:
            print(mode, "madvise(POPULATE_WRITE) errno", ctypes.get_errno())
    t1 = time.perf_counter()
    np.copyto(dst, src)
    t2 = time.perf_counter()
    np.copyto(dst, src)
    t3 = time.perf_counter()
    print("%-18s madvise %.1f ms, first copy %.1f ms (%.2f GB/s), second copy %.1f ms (%.2f GB/s)" % (
        mode, 1e3 * (t1 - t0), 1e3 * (t2 - t1), nbytes / (t2 - t1) / 1e9, 1e3 * (t3 - t2), nbytes / (t3 - t2) / 1e9))
    del dst
