"""Sharded (one MSA across N GPUs) build vs the single-GPU build, bit for bit.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/sharded_check.py [cfg2 0.1] [--time]

Every rank first builds alone (world 1), then joins the NCCL communicator and builds again
with suffix sorting sharded by k-mer key range and f(x) by column range; SA, ISA, LCP, f,
boundaries and the graph arrays must be identical.  Exit code 0 = all ranks agree."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth


def build(ctx, msa, elastic, ignore, full, index_first=True):
    ctx.set_options(elastic=elastic, gap_limit=1, ignore_chars=ignore)
    ctx.upload(msa)
    t0 = time.perf_counter()
    out = {}
    if index_first:  # explicit full index (and its read-backs); otherwise the stages build what they need, as bench.py does
        ctx.build_index()
        if full:
            out["sa"], out["isa"], out["lcp"] = ctx.sa(), ctx.isa(), ctx.lcp()
    out["f"] = ctx.compute_f()
    b, score = ctx.segment()
    g = ctx.build_graph(want_paths=False)
    torch.cuda.synchronize()
    out["secs"] = time.perf_counter() - t0
    out["b"], out["score"] = b, score
    out["nodes"], out["es"], out["ed"] = g.node_row, g.edge_src, g.edge_dst
    return out


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    cases = []
    if args:
        msa, c = synth.make_config(args[0], float(args[1]) if len(args) > 1 else 1.0)
        cases.append((args[0], msa, c["elastic"], c["ignore"]))
    else:
        rng = np.random.default_rng(3)
        cases.append(("small gapped", synth.random_small_msa(rng, 9, 120, gap_p=0.2), True, b""))
        cases.append(("tiny", np.frombuffer(b"ACAG", np.uint8).reshape(4, 1).copy(), True, b""))
        cases.append(("identical rows", np.tile(np.frombuffer(b"ACGTTGCAAC" * 30, np.uint8), (7, 1)), True, b""))
        cases.append(("one symbol", np.full((3, 3000), ord("A"), np.uint8), True, b""))
        cases.append(("msa-like gapped N", synth.make_msa(60, 5000, snp=0.002, gap_div=500, n_div=2000, seed=11), True, b"N"))
        cases.append(("cfg2 twin", synth.make_msa(1000, 3000, snp=0.001, seed=2), False, b""))
        # long inputs: many DP chunks per GPU (guess and verify are sharded) and several backtrack segments
        cases.append(("long elastic", synth.make_msa(12, 40000, snp=0.02, gap_div=300, seed=21), True, b""))
        cases.append(("long non-elastic", synth.make_msa(10, 30000, snp=0.02, seed=22), False, b""))
        # 1000+ rows with gap runs: the M list of the suffix sorter, buckets dealt out by column, repeats kept by every GPU
        cases.append(("cfg3 twin", synth.make_msa(5000, 2000, snp=0.001, gap_div=2000, seed=3), True, b""))
        rep = synth.make_msa(1500, 600, snp=0.002, gap_div=300, seed=5)
        rep[:, 300:420] = rep[:, 100:220]  # a repeat across columns (and across the GPUs' column ranges)
        cases.append(("repeat across columns", rep, True, b""))
    ctx = pkg.Context(local)
    uid = [pkg.api.dist_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ok = True
    for name, msa, elastic, ignore in cases:
        full = msa.size <= 4_000_000
        ref = build(ctx, msa, elastic, ignore, full)
        ctx.dist_init(rank, world, uid[0])
        got = build(ctx, msa, elastic, ignore, full)
        got = build(ctx, msa, elastic, ignore, full) if "--time" in sys.argv else got
        lean = build(ctx, msa, elastic, ignore, full, index_first=False)
        st = ctx.stats()
        ctx.dist_shutdown()
        # a fresh communicator per case needs a fresh id
        uid = [pkg.api.dist_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        same = all(np.array_equal(ref[k], got[k]) for k in ref if k != "secs") and ref["score"] == got["score"]
        same = same and all(np.array_equal(ref[k], lean[k]) for k in lean if k != "secs") and ref["score"] == lean["score"]
        ok = ok and same
        print("[rank %d] %-20s %s  single %.1f ms, sharded x%d %.1f ms, comm %.2f ms, rounds %d, scratch arena %.2f GB" % (
            rank, name, "identical" if same else "MISMATCH", 1e3 * ref["secs"], world, 1e3 * got["secs"],
            st["kernels"]["comm"]["ms"], st["sa_rounds"], st["arena_peak_bytes"] / 1e9), flush=True)
    flag = torch.tensor([0 if ok else 1])
    dist.all_reduce(flag)
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 0 else 1)


if __name__ == "__main__":
    main()
