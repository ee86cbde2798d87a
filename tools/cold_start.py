#!/usr/bin/env python
"""Where the first call of a process goes (the drop-in binary pays it once per run): context creation, sequence
upload, the first submit (scratch allocation) against the second (everything warm).  2 000-genome bench set."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

lz = bench.lz
ss = bench.make_set(int(sys.argv[1]) if len(sys.argv) > 1 else 2000)
n = len(ss)
t0 = time.perf_counter(); m = lz.Matcher(device=0); t1 = time.perf_counter(); print("create %.3f s" % (t1 - t0))
m.set_sequences(ss); t2 = time.perf_counter(); print("set_sequences %.3f s" % (t2 - t1))
rows = np.arange(n, dtype=np.uint32); out = np.zeros((n * (n - 1), 3), dtype=np.int32)
m.run_rows(rows, out=out); t3 = time.perf_counter(); print("first run %.3f s (device %.1f ms)" % (t3 - t2, m.timing()["total_ms"]))
m.run_rows(rows, out=out); t4 = time.perf_counter(); print("second run %.3f s (device %.1f ms)" % (t4 - t3, m.timing()["total_ms"]))
m.close(); print("close %.3f s" % (time.perf_counter() - t4))
