"""Wall-clock per stage of one build (debug aid, not a bench): python tools/stage_times.py [cfg] [scale]"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
msa, c = synth.make_config(name, scale)
m, n = msa.shape
ctx = pkg.Context(0)
ctx.set_options(elastic=c["elastic"], gap_limit=1, ignore_chars=c["ignore"])
pin = torch.from_numpy(msa).pin_memory()
dev = torch.from_numpy(msa).cuda()
def T(label, fn):
    torch.cuda.synchronize(); t = time.perf_counter(); r = fn(); torch.cuda.synchronize(); print("%-28s %8.2f ms" % (label, 1e3 * (time.perf_counter() - t))); return r
for rep in range(3):
    print("--- rep", rep, name, m, n)
    T("upload pinned host", lambda: ctx.upload_ptr(pin.data_ptr(), m, n))
    T("upload pageable host", lambda: ctx.upload(msa))
    T("upload device", lambda: ctx.upload_device(dev.data_ptr(), m, n))
    T("build_index", ctx.build_index)
    T("compute_f", lambda: ctx.lib.fbg_compute_f(ctx.h))
    T("segment", ctx.segment)
    T("build_graph", ctx.build_graph)
    T("run_device (all)", lambda: ctx.run_ptr(dev.data_ptr(), m, n, True))
    T("run_host pinned (all)", lambda: ctx.run_ptr(pin.data_ptr(), m, n, False))
st = ctx.stats()
print({k: (round(v["ms"], 2), v["launches"]) for k, v in st["kernels"].items()}, "rounds", st["sa_rounds"], "dp_passes", st["dp_passes"], "kmer_dict", st["kmer_dict"], "arena GB", st["arena_peak_bytes"] / 1e9)
