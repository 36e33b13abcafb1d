"""One sharded build of a config on all ranks (torchrun), per-kernel-group times of rank 0: debug / profiling aid."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
msa, c = synth.make_config(name, scale)
m, n = msa.shape
ctx = pkg.Context(local)
ctx.set_options(elastic=c["elastic"], gap_limit=1, ignore_chars=c["ignore"])
# the NCCL id travels through a file (no torch.distributed: rank 0 may be running under ncu and come up late)
idf = "/tmp/fbg_ncclid_%s" % os.environ.get("MASTER_PORT", "0")
if rank == 0:
    with open(idf + ".tmp", "wb") as fh:
        fh.write(pkg.api.dist_unique_id())
    os.rename(idf + ".tmp", idf)
while not os.path.exists(idf):
    time.sleep(0.05)
uid = open(idf, "rb").read()
ctx.dist_init(rank, world, uid)
dev = torch.from_numpy(msa).cuda()
for r in range(reps):
    ctx.reset_stats()
    torch.cuda.synchronize()
    t = time.perf_counter()
    nb, score, info = ctx.run_xgfa_ptr(dev.data_ptr(), m, n, device=True)
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t)
    st = ctx.stats()
    if rank == 0:
        print("rep %d: %s %dx%d on %d GPUs %.2f ms  blocks %d score %d" % (r, name, m, n, world, ms, nb, score), flush=True)
if rank == 0:
    ks = st["kernels"]
    for nm, v in sorted(ks.items(), key=lambda kv: -kv[1]["ms"]):
        if v["launches"]:
            print("  %-14s %9.3f ms  %4d launches" % (nm, v["ms"], v["launches"]))
ctx.dist_shutdown(); ctx.close()
if rank == 0:
    os.remove(idf)
