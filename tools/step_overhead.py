#!/usr/bin/env python
"""Where a resident step's time goes besides its two kernels: the same step loop as bench.py with per-kernel profiling events on / off."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from svtrek_b200 import api, simreads as S
rs, truth = S.make_config("c2")
w = S.windows_from_truth(truth)
dev = torch.device("cuda", 0)
st = torch.cuda.Stream()
a = api.Auditor(0, cuda_stream=st.cuda_stream)
a.load_reads(rs)
t = {k: torch.from_numpy(np.ascontiguousarray(w[k]).view(np.int32 if w[k].dtype.itemsize == 4 else np.uint8)).to(dev) for k in ("chrom", "start", "end", "pos", "kind")}
n = len(w["chrom"]); res = torch.empty(n, dtype=torch.int32, device=dev)
args = (n, t["chrom"].data_ptr(), t["start"].data_ptr(), t["end"].data_ptr(), t["pos"].data_ptr(), t["kind"].data_ptr(), res.data_ptr())
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
torch.cuda.synchronize()
for prof in (False, True, False, True):
    a.set_profiling(prof)
    evs = []
    with torch.cuda.stream(st):
        for it in range(40):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st); a.index_reads(); a.refine_windows_device(*args); e1.record(st); a.sync()
            if it >= 10: evs.append((e0, e1))
    torch.cuda.synchronize()
    ms = [x.elapsed_time(y) for x, y in evs]
    line = "profiling %-5s step_us mean %.1f min %.1f" % (prof, 1e3 * sum(ms) / len(ms), 1e3 * min(ms))
    if prof:
        tm = a.timing()
        line += "  index_us %.1f refine_us %.1f" % (1e3 * tm["index_kernel_ms"] / tm["index_kernel_launches"], 1e3 * tm["refine_kernel_ms"] / tm["refine_kernel_launches"])
    print(line)
