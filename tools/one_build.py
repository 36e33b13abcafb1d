"""Builds one synthetic config REPS times (MSA resident in HBM -> xGFA bytes in HBM) and prints the
per-kernel-group CUDA-event times of the last build: python tools/one_build.py cfg3 [scale] [reps]
Debug / profiling aid (run it under ncu for a launch list); not a bench."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import founderblockgraphs_b200 as pkg  # noqa: E402
from founderblockgraphs_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
msa, c = synth.make_config(name, scale)
m, n = msa.shape
ctx = pkg.Context(0)
ctx.set_options(elastic=c["elastic"], gap_limit=1, ignore_chars=c["ignore"])
dev = torch.from_numpy(msa).cuda()
for r in range(reps):
    ctx.reset_stats()
    torch.cuda.synchronize()
    t = time.perf_counter()
    nb, score, info = ctx.run_xgfa_ptr(dev.data_ptr(), m, n, device=True)
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t)
    st = ctx.stats()
    print("rep %d: %s %dx%d  %.2f ms  %.3g cells/s  blocks %d score %d xgfa %d B" % (r, name, m, n, ms, m * n / ms * 1e3, nb, score, info["n_bytes"]))
ks = st["kernels"]
tot = sum(v["ms"] for v in ks.values())
for nm, v in sorted(ks.items(), key=lambda kv: -kv[1]["ms"]):
    if v["launches"]:
        print("  %-14s %9.3f ms %5.1f %%  %4d launches  %8.1f GB/s algorithmic" % (nm, v["ms"], 100 * v["ms"] / tot, v["launches"], v["bytes"] / 1e6 / max(v["ms"], 1e-9)))
print(json.dumps(dict(total_kernel_ms=tot, wall_ms=ms, rounds=st["sa_rounds"], dp_passes=st["dp_passes"], kmer_dict=st["kmer_dict"],
                      index_depth=st["index_depth"], text_len=st["text_len"], arena_gb=st["arena_peak_bytes"] / 1e9)))
