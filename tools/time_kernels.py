#!/usr/bin/env python
"""Kernel-only timing of the two product kernels on one workload (for A/B experiments: SVTK_LIB=<variant .so>)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from svtrek_b200 import api, simreads as S
name = sys.argv[1] if len(sys.argv) > 1 else "c2"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rs, truth = S.make_config(name, scale=scale)
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
for it in range(60):
    if it == 10: a.set_profiling(True)
    with torch.cuda.stream(st): flush.zero_()
    a.index_reads(); a.refine_windows_device(*args); a.sync()
tm = a.timing()
print(os.environ.get("SVTK_LIB", "default"), name, "index_us %.1f refine_us %.1f" % (1e3 * tm["index_kernel_ms"] / tm["index_kernel_launches"], 1e3 * tm["refine_kernel_ms"] / tm["refine_kernel_launches"]), "checksum", int(res.cpu().numpy().astype(np.int64).sum()))
