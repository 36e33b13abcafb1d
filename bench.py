#!/usr/bin/env python
"""bench.py -- MSA cells/sec of the end-to-end founder-graph build (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg3]

Workload: cfg3 (BASELINE.json configs[2]: synthetic 5000 rows x 100 000 columns with indel gap
runs, --elastic): the largest BASELINE configuration that fits one GPU, and one that runs the
whole path (full SA / ISA / LCP index, per-column K2).  cfg2 (configs[1], gapless, the virtual
index) is available with --workload cfg2.

One "step" = one pass of the whole hot path (K0 pack -> K1 SA/LCP -> K2 f(x) -> K3 DP ->
K4 nodes/edges -> K5 xGFA bytes) over one synthetic MSA: from the MSA in memory to the bytes
of the output file ("after FASTA parse to xGFA bytes ready").  Two MSA variants
(different seeds) alternate between steps.
  value : cells/s with the MSA already resident in HBM when the timed region starts and the
          xGFA bytes left in HBM (CUDA events on the library's own stream, max over ranks)
  e2e   : the same through the reference-facing C-ABI calls with a pinned HOST buffer;
          H2D of the MSA and D2H of the xGFA bytes are inside the timed region
  parity_sha256_ok : once, outside the timed region, the sha256 of the xGFA bytes of both
          variants is compared with the digest of the REFERENCE's own output on the same MSA
          (tests/golden/large_cases.json, made by tests/golden/make_golden_large.py)
  roofline     : the kernel group with the largest share of device time (CUDA events around
                 every launch group): algorithmic HBM bytes / event time; ncu's DRAM bytes for
                 the same launches from profiles/dominant_kernel_traffic.json
  cpu_baseline : the reference's own CPU code (oracle/_ref/ref_harness, built from
                 /root/reference) with --threads = host cores on a bounded column sample
N > 1 (torchrun, one process per GPU): ONE MSA of the workload is built by all GPUs together
(K1: tied buckets dealt out by column range, ranks read from their owner over NVLink; K2 / DP by
column range; NCCL all-reduces at the end): `value` is that build, so the work is the same at every
N ("scaling": "strong", also on the N = 1 line); every rank checks the digest of its own result.  The old
arm -- one independent MSA per GPU, no collective -- is kept as the extra key "replicas".
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

DOMINANT_NAMES = {
    "tile_sort": "k1_tile_sort_kernel / k1_bucket_sort_kernel (bucket-local sort of the tied suffixes of one doubling round, K1)",
    "radix_scatter": "rs_scatter_kernel (radix scatter passes of K1)",
    "doubling_aux": "k1_cls_write_kernel + k1_dict_emit_tile_kernel (list compaction / round-0 emission with fused rank look-up, K1)",
    "fcol": "k2_fcol_kernel (per-column f(x))",
    "kmer": "k1_kmer_tile_kernel / k1_kmer_kernel (k-mer dictionary insert of every text position, K1 round 0)",
    "graph": "k4_hash / k4_group / k4_verify (label hash, per-block dedup, byte-wise verification, K4)",
    "edges": "k4_nodes / k4_edges / k4_edge_gather (node ids, per-block edge sort/unique, K4)",
    "pack": "k0_mask / k0_chunk_* / k0_text (bitmaps, rank tables, gap-free text, K0)",
    "dp": "k3_* (segmentation DP + parallel backtrack, K3)",
    "scan": "scan_lookback_kernel (single-pass prefix scans)",
    "lcp": "k1_irr_* (irreducible LCPs; the fill is a max-scan, K1)",
}
METRIC = "msa_cells_per_sec_end_to_end_efg_build"
UNIT = "cells/s"
GOLDEN = os.path.join(ROOT, "tests", "golden", "large_cases.json")


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic():
    """ncu evidence for the dominant kernel group (DRAM bytes per step, what bounds it)."""
    path = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
    if os.path.exists(path):
        with open(path) as fh:
            return json.load(fh)
    return None


def golden_digests(name):
    """{variant: record} of the reference digests for the full-size MSAs of a workload."""
    out = {}
    if os.path.exists(GOLDEN):
        with open(GOLDEN) as fh:
            for c in json.load(fh):
                for v in (0, 1):
                    if c["name"] in ("%s_full_v%d" % (name, v), "%s_full_nonelastic_v%d" % (name, v)):
                        out[v] = c
    return out


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons DURING the timed region through NVML (in-process:
    spawning nvidia-smi every 100 ms takes the driver lock and slows the CUDA calls being timed)."""

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            uuid = None
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            except Exception:
                pass
            h = None
            if uuid:
                for cand in ("GPU-" + uuid, uuid):
                    try:
                        h = nv.nvmlDeviceGetHandleByUUID(cand.encode() if isinstance(cand, str) else cand)
                        break
                    except Exception:
                        h = None
            if h is None:
                h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
            while not self.stop_flag:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for nm, bit in bits.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.02)
        except Exception as e:  # NVML unavailable: record that rather than guessing
            self.reasons.add("nvml_unavailable:" + type(e).__name__)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples), "how": "NVML, 20 ms period, during both timed regions"}


def workload_msa(name, rank, variant):
    from founderblockgraphs_b200 import synth
    c = dict(synth.CONFIGS[name])
    msa = synth.make_msa(c["m"], c["n"], c["snp"], c["gap_div"], c["n_div"], seed=int(name[3:]) + 100 * rank + 7 * variant)
    return msa, c


def sample_cols(name, c, requested):
    # elastic: about 5 s of reference CPU time per step (cfg3: 5000 x 1000 cells); cfg2's segment() is
    # single-threaded and takes ~100 s for the full MSA, so a fifth of it
    default = c["n"] // 100 if c["elastic"] else c["n"] // 5
    return max(64, min(c["n"], requested or default))


# --------------------------------------------------------------------- reference arm
class ReferenceSample:
    """The reference's own CPU implementation (oracle/_ref, built from /root/reference) on the first
    `cols` columns of variant 0 of the workload; the FASTA file is written once."""

    def __init__(self, name, cols):
        from oracle import oracle as O
        self.O = O
        msa, c = workload_msa(name, 0, 0)
        self.c = c
        self.msa = np.ascontiguousarray(msa[:, :cols])
        self.path = "/tmp/fbg_bench_%d.fasta" % os.getpid()
        O.write_fasta(self.path, [bytes(r) for r in self.msa])
        self.cores = os.cpu_count() or 1
        self.sample = "first %d of %d columns x %d rows of %s variant 0 (%s)" % (cols, c["n"], c["m"], name, "--elastic --gfa, --threads=%d" % self.cores if c["elastic"] else "non-elastic gap-limit=1")

    def run(self):
        O, c = self.O, self.c
        if O.have_reference():
            kind = "reference"
            if c["elastic"]:
                args = ["--threads", str(self.cores), "--no-dump"] + (["--ignore-chars", c["ignore"].decode()] if c["ignore"] else [])
                res, _, _ = O.run_harness(self.path, "elastic", args, timeout=3000, want_output=True)
                used = self.cores  # compute_f is threaded (fbg.cpp:2278-2289); index construction and the DP are serial
            else:
                res, _, _ = O.run_harness(self.path, "nonelastic", ["--gap-limit", "1", "--no-dump"], timeout=3000)
                used = 1  # segment() is single-threaded by construction (fbg.cpp:526-735)
            secs = res["time_index_s"] + res["time_segment_s"] + res.get("time_write_s", 0.0)
        else:
            kind, used = "port", 1
            t0 = time.time()
            f = O.compute_f(self.msa, c["ignore"], not c["elastic"])
            b = O.dp_elastic(f)[2] if c["elastic"] else O.dp_nonelastic(O.v_from_f(f))[3]
            O.graph(self.msa, b)
            secs = time.time() - t0
        cells = self.msa.shape[0] * self.msa.shape[1]
        return cells / secs, secs, kind, used

    def close(self):
        if os.path.exists(self.path):
            os.remove(self.path)


def full_size_reference(name):
    """The reference on the FULL workload, measured once in the build container (the record
    tests/golden/make_golden_large.py leaves; also quoted in BASELINE.md)."""
    g = golden_digests(name).get(0)
    if not g:
        return None
    secs = (g["ref_time_index_s"] or 0) + (g["ref_time_segment_s"] or 0) + (g["ref_time_write_s"] or 0)
    return {"cells_per_s": g["m"] * g["n"] / secs, "seconds": secs, "index_s": g["ref_time_index_s"], "segment_s": g["ref_time_segment_s"],
            "write_s": g["ref_time_write_s"], "threads": g["threads"], "host": g["ref_host"], "where": "tests/golden/large_cases.json:" + g["name"]}


def config_dict(name, c, extra=None):
    d = {"workload": "%s: synthetic %d rows x %d columns, SNP %.3g%s, %s%s" % (
        name, c["m"], c["n"], c["snp"], ", indel gap runs" if c["gap_div"] else ", gapless",
        "--elastic --gfa" if c["elastic"] else "non-elastic --gap-limit=1",
        (" --ignore-chars=" + c["ignore"].decode()) if c["ignore"] else ""),
        "rows": c["m"], "columns": c["n"], "cells_per_step": c["m"] * c["n"],
        "l2": "the MSA alone (%d MB) exceeds the 126 MB L2; a 256 MiB buffer is also rewritten between timed steps" % (c["m"] * c["n"] // 1000000)}
    if extra:
        d.update(extra)
    return d


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from founderblockgraphs_b200 import synth
    c = dict(synth.CONFIGS[args.workload])
    cols = sample_cols(args.workload, c, args.sample_cols)
    rs = ReferenceSample(args.workload, cols)
    vals, secs_all = [], []
    kind = used = None
    try:
        for it in range(args.warmup + args.steps):
            v, secs, kind, used = rs.run()
            if it >= args.warmup:
                vals.append(v)
                secs_all.append(secs)
    finally:
        rs.close()
    value = float(np.mean(vals))
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * float(np.mean(secs_all)), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u8/u32", "data": "synthetic", "config": config_dict(args.workload, c, {"same_config": cols == c["n"]}),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": kind, "sample": rs.sample, "host_cores": os.cpu_count(),
                             "full_size": full_size_reference(args.workload)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def sha256_of_xgfa(ctx, nbytes):
    h = hashlib.sha256()
    step = 64 << 20
    for off in range(0, nbytes, step):
        h.update(ctx.read_xgfa(off, min(step, nbytes - off)))
    return h.hexdigest()


def run_ours(args):
    # rank 0 prints ONE JSON line on stdout: everything else that writes to file descriptor 1 (NCCL's
    # "NCCL version ..." banner when NCCL_DEBUG is set, library chatter) is sent to stderr, and the
    # line goes out through a saved copy of the original descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    import founderblockgraphs_b200 as pkg

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    name = args.workload
    c = dict(pkg.synth.CONFIGS[name])
    # N = 1: this GPU's MSAs.  N > 1: every rank holds rank 0's MSAs (ONE MSA built by all GPUs together)
    variants = [workload_msa(name, 0, v)[0] for v in range(2)]
    m, n = variants[0].shape
    cells = m * n
    ctx = pkg.Context(local)
    ctx.set_options(elastic=c["elastic"], gap_limit=1, ignore_chars=c["ignore"])
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(pkg.api.dist_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        ctx.dist_init(rank, world, bytes(uid.cpu().numpy().tobytes()))
    ext = torch.cuda.ExternalStream(ctx.stream_ptr(), device=torch.device("cuda", local))
    dev_msa = [torch.from_numpy(v).cuda() for v in variants]
    pin_msa = [torch.from_numpy(v).pin_memory() for v in variants]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity, outside every timed region: digest of the xGFA of both variants against the reference's
    gold = golden_digests(name)
    parity = {"checked_variants": [], "golden": "tests/golden/large_cases.json (reference run by tests/golden/make_golden_large.py)"}
    ok_all = bool(gold)
    info0 = None
    for v in range(2):
        nb, score, info = ctx.run_xgfa_ptr(dev_msa[v].data_ptr(), m, n, device=True)
        info0 = info0 or info
        if v in gold:
            g = gold[v]
            good = (sha256_of_xgfa(ctx, info["n_bytes"]) == g["xgfa_sha256"] and info["n_bytes"] == g["xgfa_bytes"]
                    and nb == g["n_blocks"] and score == g["score"])
            parity["checked_variants"].append({"variant": v, "case": g["name"], "ok": bool(good), "xgfa_bytes": int(info["n_bytes"]), "blocks": int(nb), "score": int(score)})
            ok_all = ok_all and good
    if world > 1:
        t_ok = torch.tensor([1 if ok_all else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        ok_all = bool(int(t_ok.item()))
        parity["ranks_checked"] = world
    pin_out = torch.empty(int(info0["n_bytes"] * 1.25) + 4096, dtype=torch.uint8).pin_memory()
    last_info = [info0]

    def one_step(i, host):
        src = pin_msa[i % 2] if host else dev_msa[i % 2]
        flush.zero_()
        barrier() if world > 1 else torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(ext)
        want_out = host and rank == 0
        nb, score, last_info[0] = ctx.run_xgfa_ptr(src.data_ptr(), m, n, device=not host,
                                                    host_out=pin_out.data_ptr() if want_out else 0, host_cap=pin_out.numel())
        b.record(ext)
        b.synchronize()
        wall = time.perf_counter() - t0
        return a.elapsed_time(b) / 1000.0, wall, nb, score

    # ---- device-resident arm
    for i in range(args.warmup):
        one_step(i, False)
    ctx.reset_stats()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    dev_secs = [one_step(i, False)[0] for i in range(args.steps)]
    barrier()
    st = ctx.stats()
    # ---- end-to-end arm (host buffers through the C ABI)
    for i in range(min(args.warmup, 2)):
        one_step(i, True)
    barrier()
    e2e = [one_step(i, True) for i in range(args.steps)]
    barrier()
    sampler.stop_flag = True
    sampler.join()
    d2h = int(last_info[0]["n_bytes"])  # the xGFA bytes of the step (rank 0 reads them back)
    nb, score = e2e[-1][2], e2e[-1][3]

    # sharded build: a step ends when the slowest rank ends (the steps are separated by barriers): max over ranks per step
    t_steps = torch.tensor([dev_secs, [x[1] for x in e2e]], dtype=torch.float64, device="cuda")
    per_rank_ms = [1000.0 * float(np.mean(dev_secs))]
    if world > 1:
        gathered = [torch.zeros_like(t_steps) for _ in range(world)]
        dist.all_gather(gathered, t_steps)
        per_rank_ms = [1000.0 * float(g[0].mean().item()) for g in gathered]
        dist.all_reduce(t_steps, op=dist.ReduceOp.MAX)
    t_dev_total, t_e2e_total = float(t_steps[0].sum().item()), float(t_steps[1].sum().item())
    total_cells = cells * args.steps  # ONE MSA per step, whatever the number of GPUs
    value = total_cells / t_dev_total
    e2e_value = total_cells / t_e2e_total

    # ---- N > 1, extra: the old arm, one independent MSA per GPU and no collective (weak scaling)
    replicas = None
    if world > 1 and not args.no_replicas:
        ctx.dist_shutdown()
        own = [torch.from_numpy(workload_msa(name, rank, v)[0]).cuda() for v in range(2)] if rank else dev_msa
        rsteps = max(2, min(args.steps, 5))

        def rep_step(i):
            flush.zero_()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(ext)
            ctx.run_xgfa_ptr(own[i % 2].data_ptr(), m, n, device=True)
            b.record(ext)
            b.synchronize()
            return a.elapsed_time(b) / 1000.0
        for i in range(2):
            rep_step(i)
        barrier()
        rsecs = [rep_step(i) for i in range(rsteps)]
        barrier()
        t_rep = torch.tensor([sum(rsecs)], dtype=torch.float64, device="cuda")
        each = [torch.zeros_like(t_rep) for _ in range(world)]
        dist.all_gather(each, t_rep)
        dist.all_reduce(t_rep, op=dist.ReduceOp.MAX)
        replicas = {"what": "one independent MSA of the workload per GPU, no data-path collective", "scaling": "weak", "steps": rsteps,
                    "value": cells * world * rsteps / float(t_rep.item()), "unit": UNIT, "ms_per_step": 1000.0 * float(t_rep.item()) / rsteps,
                    "per_rank_ms": [round(1000.0 * float(e.item()) / rsteps, 3) for e in each]}

    if rank == 0:
        peak, peak_src = load_peaks()
        ks = st["kernels"]
        compute = {nm: v for nm, v in ks.items() if nm != "comm"}
        dom = max(compute, key=lambda nm: compute[nm]["ms"])  # the kernel group with the largest share of device time
        k = ks[dom]
        achieved = (k["bytes"] / 1e9) / (k["ms"] / 1e3) if k["ms"] > 0 else 0.0
        traffic = load_traffic()
        t_match = traffic if traffic and traffic.get("group") == dom and traffic.get("workload") == name else None
        total_ms = sum(v["ms"] for v in ks.values())
        launches = int(sum(v["launches"] for v in ks.values()))
        cols = sample_cols(name, c, args.sample_cols)
        rs = ReferenceSample(name, cols)
        try:
            cb_v, cb_secs, cb_kind, cb_cores = rs.run()
        finally:
            rs.close()
        roof = {"bound": t_match["bound"] if t_match else "hbm", "kernel": DOMINANT_NAMES.get(dom, dom), "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": k["bytes"] / max(1, k["launches"]), "launches": int(k["launches"]), "launches_per_step": k["launches"] / args.steps,
                "ms_per_launch": k["ms"] / max(1, k["launches"]),
                "what_counts": "algorithmic bytes = compulsory HBM traffic of the kernel as written (DESIGN.md section 3); tables that stay in L2 are not counted",
                "traffic": (t_match["dram_bytes_per_step"] / t_match["launches_per_step"]) if t_match else None,
                "share_of_kernel_time": k["ms"] / total_ms if total_ms else None,
                "other_kernels_hbm_frac": {nm: round(((v["bytes"] / 1e9) / (v["ms"] / 1e3)) / peak, 4) for nm, v in ks.items() if v["ms"] > 0 and nm not in (dom, "comm")},
                # all algorithmic bytes of a step over the step's device time
                "pipeline": {"algorithmic_bytes_per_step": sum(v["bytes"] for nm, v in ks.items() if nm != "comm") / args.steps,
                             "frac": ((sum(v["bytes"] for nm, v in ks.items() if nm != "comm") / 1e9) / (t_dev_total if world == 1 else total_ms / 1e3)) / peak if t_dev_total > 0 else 0.0}}
        if t_match:
            roof["traffic_evidence"] = t_match.get("evidence")
            roof["frac_from_traffic"] = (t_match["dram_bytes_per_step"] / 1e9) / ((k["ms"] / args.steps) / 1e3) / peak if k["ms"] > 0 else None
            roof["note"] = t_match.get("note", "")
        else:
            roof["note"] = "no ncu capture of this kernel group on this workload is committed under profiles/ (traffic null, bound assumed)"
        par = ("ONE MSA built by all %d GPUs: K1 sharded by k-mer key range, K2 / DP by column range, NCCL collectives over NVLink; results identical on every rank" % world) if world > 1 else "single GPU"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * t_dev_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u8/u32", "data": "synthetic",
            "parity_sha256_ok": ok_all, "parity": parity,
            "per_rank_ms": [round(x, 3) for x in per_rank_ms],
            "config": config_dict(name, c, {"blocks": int(nb), "score": int(score), "text_len": int(st["text_len"]), "sa_rounds": int(st["sa_rounds"]),
                                            "index": ("sufficient index: K1 stopped at depth %d (all tied buckets column-pure)%s" % (
                                                st["index_depth"], ", f(x) from the sorted k-mer list, no SA / ISA / LCP array" if st["index_virtual"] else "")) if st["index_depth"] else "full SA / ISA / LCP",
                                            "index_depth": int(st["index_depth"]), "index_virtual": int(st["index_virtual"]),
                                            "step": "MSA -> xGFA bytes (K0..K5)", "xgfa_bytes": int(last_info[0]["n_bytes"]),
                                            "nodes": int(last_info[0]["n_nodes"]), "edges": int(last_info[0]["n_edges"]),
                                            "parallelism": par}),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(cells) * world, "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1000.0 * t_e2e_total / args.steps,
                    "note": "every rank copies the MSA from its own pinned host buffer; rank 0 reads the xGFA bytes back" if world > 1 else "pinned host MSA in, xGFA bytes out to pinned host memory"},
            "gpu_launches": launches,
            "roofline": roof,
            "kernels": {nm: {"ms_per_step": v["ms"] / args.steps, "GBps": ((v["bytes"] / 1e9) / (v["ms"] / 1e3)) if v["ms"] > 0 else 0.0,
                             "launches_per_step": v["launches"] / args.steps} for nm, v in ks.items()},
            "cpu_baseline": {"value": cb_v, "unit": UNIT, "cores": cb_cores, "kind": cb_kind, "sample": rs.sample, "seconds": cb_secs,
                             "host_cores": os.cpu_count(), "full_size": full_size_reference(name)},
            "clocks": sampler.summary(),
        }
        if world > 1:
            line["comm_ms_per_step"] = ks["comm"]["ms"] / args.steps
            line["device_bytes_per_gpu"] = int(st["arena_peak_bytes"])
        if replicas:
            line["replicas"] = replicas
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1 and ctx.h:
        try:
            ctx.dist_shutdown()
        except Exception:
            pass
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=["cfg2", "cfg3"])
    ap.add_argument("--sample-cols", type=int, default=0, help="columns of the CPU-baseline sample (default: n/100 elastic, n/5 non-elastic)")
    ap.add_argument("--no-replicas", action="store_true", help="N > 1: skip the extra one-MSA-per-GPU arm")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
