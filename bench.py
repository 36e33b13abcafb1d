#!/usr/bin/env python
"""bench.py -- MSA cells/sec of the end-to-end founder-graph build (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]

One "step" = one pass of the whole hot path (K0 pack -> K1 SA/LCP -> K2 f(x) -> K3 DP ->
K4 nodes/edges -> K5 xGFA bytes) over one synthetic MSA of the named BASELINE shape: from the
MSA in memory to the bytes of the output file ("after FASTA parse to xGFA bytes ready",
SURVEY M1).
  value : cells/s with the MSA already resident in HBM when the timed region starts and the
          xGFA bytes left in HBM (CUDA events on the library's own stream, max over ranks)
  e2e   : the same through the reference-facing C-ABI calls with a pinned HOST buffer;
          H2D of the MSA and D2H of the xGFA bytes are inside the timed region
  roofline     : dominant kernel (radix scatter passes) algorithmic bytes / event time
  cpu_baseline : the reference's own CPU code (oracle/_ref/ref_harness, built from
                 /root/reference) on a bounded sample of the same workload
N > 1: one process per GPU (torchrun); every rank builds the graph of its own
independent MSA of the same shape (per-chromosome sharding, no data-path collective),
"scaling": "weak".  See DESIGN.md section "Multi-GPU".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

DOMINANT_NAMES = {
    "tile_sort": "k1_tile_sort_kernel (bucket-local shared-memory sort of tied suffixes, K1)",
    "radix_scatter": "rs_scatter_kernel (radix scatter passes of K1)",
    "doubling_aux": "k1_cls_write_kernel + k1_dict_emit_kernel (list compaction / round-0 emission with fused rank look-up, K1)",
    "fcol": "k2_fcol_kernel (per-column f(x))",
    "kmer": "k1_kmer_kernel<5> (virtual index: every text position looks its k-mer up in / enters it into the open-addressing table "
            "and checks the slot's column, K1 round 0; <4>, the full insert, when per-position ranks are needed)",
    "graph": "k4_hash / k4_group / k4_verify (label hash, per-block dedup, byte-wise verification, K4)",
    "edges": "k4_nodes / k4_edges / k4_edge_gather (node ids, per-block edge sort/unique, K4)",
    "pack": "k0_mask / k0_chunk_* / k0_text (bitmaps, rank tables, gap-free text, K0)",
    "dp": "k3_* (chunked guess-and-verify segmentation DP + parallel backtrack, K3)",
    "scan": "scan_lookback_kernel (single-pass prefix scans)",
}
ROOFLINE_NOTES = {
    "graph": "a group of six short kernels (label hash from the text, per-block shared-memory sort of (hash, row), byte-wise verification, "
             "node ids, per-block edge sort/unique, gather): algorithmic bytes 2 B per text character + 60 B per (row, block) record; at "
             "cfg2 (2.6 M records) every one of them runs for tens of microseconds, i.e. at launch / latency scale, not at HBM scale",
    "pack": "algorithmic bytes 2 B per cell + 5 B per text character + 24 B per (row, 32-column tile)",
    "kmer": "algorithmic bytes 17 B per text character (1 B text, 8 B slot + arrival number written, 8 B for the two L2 atomics' operands); "
            "with the virtual index (gapless MSA, cfg2) the light form <5> runs: 13 B per text character (1 B text, the 8 B key word and the "
            "4 B column word of its table slot, requested from L2); ncu: 79 MB of DRAM traffic per launch against 390 MB requested - the 50 MB "
            "table stays in the 126 MB L2, only the text streams from HBM - DRAM 2 % of peak, L2 41 %, issue slots 22 % busy, 37 % achieved "
            "occupancy (58 registers): the kernel waits for random L2 sectors of the 4 M-slot table (long-scoreboard stalls), not for HBM; "
            "the full form <4> adds three L2 atomics per position",
    "tile_sort": "algorithmic HBM bytes are 25 B per tied suffix per round (read lp,lr,r2; write lp,lr,cls; 4 B rank scatter); ncu DRAM "
                 "traffic is 1.05x that (no re-reads).  The kernel is bound by instruction issue (61-72 % of the issue slots busy: head "
                 "scans, prefix counts and the shared-memory sorting network of each bucket's minority), not by HBM "
                 "(profiles/README.md) - it replaced seven device-wide radix passes (470 B per suffix per round)",
}
METRIC = "msa_cells_per_sec_end_to_end_efg_build"
UNIT = "cells/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture."""
    path = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
    if os.path.exists(path):
        with open(path) as fh:
            return json.load(fh)
    return None


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


# --------------------------------------------------------------------- reference arm
def reference_sample(name, cols):
    """Times the reference's own CPU implementation on the first `cols` columns."""
    from oracle import oracle as O
    msa, c = workload_msa(name, 0, 0)
    msa = np.ascontiguousarray(msa[:, :cols])
    tmp = "/tmp/fbg_bench_%d.fasta" % os.getpid()
    O.write_fasta(tmp, [bytes(r) for r in msa])
    cores = os.cpu_count() or 1
    try:
        if O.have_reference():
            kind = "reference"
            if c["elastic"]:
                args = ["--threads", str(cores), "--no-dump"] + (["--ignore-chars", c["ignore"].decode()] if c["ignore"] else [])
                res, err, _ = O.run_harness(tmp, "elastic", args, timeout=3000, want_output=True)
                used = cores
            else:
                res, err, _ = O.run_harness(tmp, "nonelastic", ["--gap-limit", "1", "--no-dump"], timeout=3000)
                used = 1  # segment() is single-threaded by construction (fbg.cpp:526-735)
            secs = res["time_index_s"] + res["time_segment_s"] + res.get("time_write_s", 0.0)
        else:
            kind, used = "port", 1
            t0 = time.time()
            f = O.compute_f(msa, c["ignore"], not c["elastic"])
            if c["elastic"]:
                _, _, b = O.dp_elastic(f)
            else:
                _, _, _, b, _ = O.dp_nonelastic(O.v_from_f(f))
            O.graph(msa, b)
            secs = time.time() - t0
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)
    cells = msa.shape[0] * msa.shape[1]
    sample = "first %d of %d columns x %d rows of %s (%s)" % (cols, c["n"], c["m"], name, "elastic" if c["elastic"] else "non-elastic gap-limit=1")
    return cells / secs, secs, kind, used, sample


def config_dict(name, c, extra=None):
    d = {"workload": "%s: synthetic %d rows x %d columns, SNP %.3g, %s%s" % (
        name, c["m"], c["n"], c["snp"], "--elastic --gfa" if c["elastic"] else "non-elastic --gap-limit=1",
        (" --ignore-chars=" + c["ignore"].decode()) if c["ignore"] else ""),
        "rows": c["m"], "columns": c["n"], "cells_per_step_per_gpu": c["m"] * c["n"],
        "l2": "working set per step (SA/ISA/LCP/sort buffers, > 1 GB) exceeds the 126 MB L2; a 256 MiB buffer is also rewritten between timed steps"}
    if extra:
        d.update(extra)
    return d


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from founderblockgraphs_b200 import synth
    c = dict(synth.CONFIGS[args.workload])
    cols = max(64, min(c["n"], args.sample_cols or c["n"] // 5))
    vals, secs_all = [], []
    kind = used = sample = None
    for it in range(args.warmup + args.steps):
        v, secs, kind, used, sample = reference_sample(args.workload, cols)
        if it >= args.warmup:
            vals.append(v)
            secs_all.append(secs)
    value = float(np.mean(vals))
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * float(np.mean(secs_all)), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/u32", "data": "synthetic", "config": config_dict(args.workload, c),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


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
    variants = [workload_msa(name, rank, v)[0] for v in range(2)]
    c = dict(pkg.synth.CONFIGS[name])
    m, n = variants[0].shape
    cells = m * n
    ctx = pkg.Context(local)
    ctx.set_options(elastic=c["elastic"], gap_limit=1, ignore_chars=c["ignore"])
    ext = torch.cuda.ExternalStream(ctx.stream_ptr(), device=torch.device("cuda", local))
    dev_msa = [torch.from_numpy(v).cuda() for v in variants]
    pin_msa = [torch.from_numpy(v).pin_memory() for v in variants]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    # host destination of the xGFA bytes in the end-to-end arm (sized by a first untimed build)
    _, _, info0 = ctx.run_xgfa_ptr(dev_msa[0].data_ptr(), m, n, device=True)
    pin_out = torch.empty(int(info0["n_bytes"] * 1.5) + 4096, dtype=torch.uint8).pin_memory()
    last_info = [info0]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(i, host):
        src = pin_msa[i % 2] if host else dev_msa[i % 2]
        flush.zero_()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(ext)
        nb, score, last_info[0] = ctx.run_xgfa_ptr(src.data_ptr(), m, n, device=not host,
                                                    host_out=pin_out.data_ptr() if host else 0, host_cap=pin_out.numel())
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
    d2h = int(last_info[0]["n_bytes"])  # the xGFA bytes of the step
    nb, score = e2e[-1][2], e2e[-1][3]

    t_dev = torch.tensor([sum(dev_secs), sum(x[1] for x in e2e)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    t_dev_total, t_e2e_total = t_dev.tolist()
    total_cells = cells * world * args.steps
    value = total_cells / t_dev_total
    e2e_value = total_cells / t_e2e_total

    # ---- N > 1: the same workload as ONE MSA sharded across all GPUs (strong scaling, NCCL collectives)
    sharded = None
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(pkg.api.dist_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        ctx.dist_init(rank, world, bytes(uid.cpu().numpy().tobytes()))
        shared = [torch.from_numpy(workload_msa(name, 0, v)[0]).cuda() for v in range(2)]  # every rank: rank 0's MSAs

        def sharded_step(i):
            flush.zero_()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(ext)
            nb_s, score_s, _ = ctx.run_xgfa_ptr(shared[i % 2].data_ptr(), m, n, device=True)
            b.record(ext)
            b.synchronize()
            return a.elapsed_time(b) / 1000.0, nb_s, score_s

        for i in range(max(2, args.warmup)):
            sharded_step(i)
        ctx.reset_stats()
        res = [sharded_step(i) for i in range(args.steps)]
        st_sh = ctx.stats()
        t_sh = torch.tensor([sum(x[0] for x in res)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t_sh, op=dist.ReduceOp.MAX)
        sharded = {"what": "ONE MSA of the workload built by all %d GPUs together: K1 sharded by k-mer key range, K2 by column range, "
                           "NCCL all-reduce(max) of rank replicas and f; results identical on every rank" % world,
                   "scaling": "strong", "value": cells * args.steps / float(t_sh), "unit": UNIT, "ms_per_step": 1000.0 * float(t_sh) / args.steps,
                   "blocks": int(res[-1][1]), "score": int(res[-1][2]),
                   "comm_ms_per_step": st_sh["kernels"]["comm"]["ms"] / args.steps,
                   "kernels_ms_per_step": {nm: round(v["ms"] / args.steps, 3) for nm, v in st_sh["kernels"].items()}}
        ctx.dist_shutdown()

    if rank == 0:
        peak, peak_src = load_peaks()
        ks = st["kernels"]
        dom = max(ks, key=lambda nm: ks[nm]["ms"])  # the kernel group with the largest share of device time
        k = ks[dom]
        achieved = (k["bytes"] / 1e9) / (k["ms"] / 1e3) if k["ms"] > 0 else 0.0
        traffic = load_traffic()
        total_ms = sum(v["ms"] for v in ks.values())
        launches = int(sum(v["launches"] for v in ks.values()))
        cols = max(64, min(c["n"], args.sample_cols or c["n"] // 5))
        cb_v, cb_secs, cb_kind, cb_cores, cb_sample = reference_sample(name, cols)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * t_dev_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/u32", "data": "synthetic",
            "config": config_dict(name, c, {"blocks": int(nb), "score": int(score), "text_len": int(st["text_len"]), "sa_rounds": int(st["sa_rounds"]),
                                            "index": ("sufficient index: K1 stopped at depth %d (all tied buckets column-pure)%s" % (
                                                st["index_depth"], ", f(x) from the sorted k-mer list" if st["index_virtual"] else "")) if st["index_depth"] else "full SA / ISA / LCP",
                                            "step": "MSA -> xGFA bytes (K0..K5)", "xgfa_bytes": int(last_info[0]["n_bytes"]),
                                            "nodes": int(last_info[0]["n_nodes"]), "edges": int(last_info[0]["n_edges"]),
                                            "parallelism": "1 MSA per GPU, independent" if world > 1 else "single GPU"}),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(cells), "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1000.0 * t_e2e_total / args.steps},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": DOMINANT_NAMES.get(dom, dom), "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "peak_source": peak_src,
                         "traffic": traffic.get("dram_bytes_per_launch") if traffic and traffic.get("group") == dom else None,
                         "note": ROOFLINE_NOTES.get(dom, ""),
                         "other_hbm_kernels": {nm: round(((v["bytes"] / 1e9) / (v["ms"] / 1e3)) / peak, 4) for nm, v in ks.items() if v["ms"] > 0 and nm != dom},
                         "algorithmic_bytes_per_launch": k["bytes"] / max(1, k["launches"]), "launches": int(k["launches"]),
                         "share_of_kernel_time": k["ms"] / total_ms if total_ms else None,
                         # all algorithmic bytes of a step over the step's device time
                         "pipeline": {"algorithmic_bytes_per_step": sum(v["bytes"] for v in ks.values()) / args.steps,
                                      "achieved": (sum(v["bytes"] for v in ks.values()) / 1e9) / t_dev_total if t_dev_total > 0 else 0.0,
                                      "frac": ((sum(v["bytes"] for v in ks.values()) / 1e9) / t_dev_total) / peak if t_dev_total > 0 else 0.0}},
            "kernels": {nm: {"ms_per_step": v["ms"] / args.steps, "GBps": ((v["bytes"] / 1e9) / (v["ms"] / 1e3)) if v["ms"] > 0 else 0.0,
                             "launches_per_step": v["launches"] / args.steps} for nm, v in ks.items()},
            "cpu_baseline": {"value": cb_v, "unit": UNIT, "cores": cb_cores, "kind": cb_kind, "sample": cb_sample, "seconds": cb_secs,
                             "host_cores": os.cpu_count()},
            "clocks": sampler.summary(),
        }
        if sharded:
            line["sharded_single_msa"] = sharded
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--sample-cols", type=int, default=0, help="columns of the CPU-baseline sample (default n/5)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
