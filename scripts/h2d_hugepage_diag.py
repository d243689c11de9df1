"""GPU-box diagnostic: is the bimodal 2 MB H2D time (45 us vs ~150 us per process) an IOMMU/page-size effect?
Compares a torch pinned buffer with a 2 MiB-aligned MADV_HUGEPAGE buffer registered with cudaHostRegister."""
import ctypes as C, mmap, os, sys, time
import numpy as np
import torch

print("THP enabled:", open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip(),
      "| defrag:", open("/sys/kernel/mm/transparent_hugepage/defrag").read().strip())
n = 131072
src = np.random.default_rng(0).standard_normal((n, 4)).astype(np.float32)
d = torch.empty((n, 4), dtype=torch.float32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(host_tensor, reps=60):
    tt = []
    for _ in range(reps):
        flush.zero_()
        torch.cuda.synchronize()
        e0.record()
        d.copy_(host_tensor, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        tt.append(e0.elapsed_time(e1) * 1e3)
    return np.median(tt), np.percentile(tt, 10), np.percentile(tt, 90)


def huge_info(addr, size):
    tot = 0
    cur = False
    for line in open("/proc/self/smaps"):
        f = line.split()
        if "-" in f[0] and len(f) >= 5 and all(c in "0123456789abcdef-" for c in f[0]):
            a, b = (int(x, 16) for x in f[0].split("-"))
            cur = a < addr + size and b > addr
        elif cur and line.startswith("AnonHugePages:"):
            tot += int(f[1])
    return tot


pinned = torch.empty((n, 4), dtype=torch.float32).pin_memory()
pinned.copy_(torch.from_numpy(src))
print("torch pinned     : H2D 2 MB median/p10/p90 us", timed(pinned), "AnonHugePages kB", huge_info(pinned.data_ptr(), n * 16))

HP = 2 << 20
mm = mmap.mmap(-1, 3 * HP, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
base = C.addressof(C.c_char.from_buffer(mm))
al = (base + HP - 1) & ~(HP - 1)
libc = C.CDLL("libc.so.6", use_errno=True)
libc.madvise.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
print("madvise(MADV_HUGEPAGE) ->", libc.madvise(al, HP, 14))
buf = (C.c_float * (n * 4)).from_address(al)
arr = np.frombuffer(buf, dtype=np.float32).reshape(n, 4)
arr[:] = src
rt = torch.cuda.cudart()
rc = rt.cudaHostRegister(al, HP, 0)
print("cudaHostRegister ->", rc, "AnonHugePages kB", huge_info(al, HP))
ht = torch.from_numpy(arr)
print("hugepage + reg   : H2D 2 MB median/p10/p90 us", timed(ht))
print("torch pinned again:", timed(pinned))
