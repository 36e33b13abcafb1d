import sys, os, time
sys.path.insert(0, "/root/repo")
import torch, numpy as np
import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth
msa, c = synth.make_config("cfg3", 1.0)
ctx = pkg.Context(0)
ctx.set_options(elastic=True)
ctx.upload(msa)
ctx.lib.fbg_compute_f(ctx.h)
b0, s0 = ctx.segment()
for chunk, warm in [(0, 0), (8192, 8192), (12288, 12288), (16384, 16384), (8192, 4096), (24576, 12288), (6144, 6144), (4096, 4096)]:
    ctx.set_dp_chunking(chunk, warm)
    ctx.reset_stats()
    torch.cuda.synchronize(); t = time.perf_counter()
    b, s = ctx.segment()
    torch.cuda.synchronize(); ms = 1e3 * (time.perf_counter() - t)
    st = ctx.stats()
    print("chunk %6d warm %6d: wall %.2f ms, dp group %.2f ms (%d launches), passes %d, same %s" % (chunk, warm, ms, st["kernels"]["dp"]["ms"], st["kernels"]["dp"]["launches"], st["dp_passes"], np.array_equal(b, b0) and s == s0), flush=True)
