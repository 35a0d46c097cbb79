#!/usr/bin/env python
"""Developer probe: ways of getting 272 MB of features into a new file on tmpfs (the ark of the bench's CLI run)."""
import mmap
import os
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

NB = 272 * 1000 * 1000
T = int(sys.argv[1]) if len(sys.argv) > 1 else 8
src = np.random.default_rng(0).integers(0, 255, NB, dtype=np.uint8)
pool = ThreadPoolExecutor(T)
d = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()


def fresh():
    p = os.path.join(d, "probe_%d.bin" % os.getpid())
    if os.path.exists(p):
        os.unlink(p)
    return p, os.open(p, os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)


def par_copy(fd, n=T):
    mm = mmap.mmap(fd, NB, access=mmap.ACCESS_WRITE)
    dst = np.frombuffer(mm, dtype=np.uint8)
    k = (NB + n - 1) // n
    list(pool.map(lambda i: dst.__setitem__(slice(i * k, min(NB, (i + 1) * k)), src[i * k:min(NB, (i + 1) * k)]), range(n)))
    del dst
    mm.close()


def timed(name, fn):
    t = time.perf_counter()
    fn()
    print("%-58s %7.1f ms" % (name, (time.perf_counter() - t) * 1e3), flush=True)


for rep in range(2):
    p, fd = fresh()
    timed("ftruncate + parallel mmap copy (%d threads), fresh pages" % T, lambda: (os.ftruncate(fd, NB), par_copy(fd)))
    timed("  same file again (pages exist)", lambda: par_copy(fd))
    os.close(fd)
    p, fd = fresh()
    timed("posix_fallocate, one call", lambda: os.posix_fallocate(fd, 0, NB))
    timed("  parallel mmap copy after fallocate", lambda: par_copy(fd))
    timed("  pwrite after fallocate, one thread", lambda: os.pwrite(fd, memoryview(src), 0))
    os.close(fd)
    p, fd = fresh()
    k = (NB + T - 1) // T
    timed("posix_fallocate in %d threads" % T, lambda: list(pool.map(lambda i: os.posix_fallocate(fd, i * k, min(k, NB - i * k)), range(T))))
    os.close(fd)
    p, fd = fresh()
    timed("pwrite fresh, one thread", lambda: os.pwrite(fd, memoryview(src), 0))
    os.close(fd)
    p, fd = fresh()
    os.ftruncate(fd, NB)
    mm = mmap.mmap(fd, NB, access=mmap.ACCESS_WRITE)
    timed("madvise(MADV_POPULATE_WRITE) one call", lambda: mm.madvise(23, 0, NB))
    mm.close()
    timed("  parallel mmap copy after populate", lambda: par_copy(fd))
    os.close(fd)
    os.unlink(p)
print(open("/sys/kernel/mm/transparent_hugepage/shmem_enabled").read().strip() if os.path.exists("/sys/kernel/mm/transparent_hugepage/shmem_enabled") else "no shmem_enabled")
os.system("df -h /dev/shm /tmp | head -5; uname -r")
