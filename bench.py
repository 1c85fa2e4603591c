#!/usr/bin/env python
"""bench.py — scored vectors/s of the encoded-vector scan on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload u8|bin|pq|u8batch] [--impl reference]

A step = one full scan of the rank's shard against one encoded query, fused top-k included, plus (N > 1) the exchange
of the per-rank top-k keys and the final merge.  The headline record is WEAK scaling: every rank holds `--rows` points
(default = the workload's BASELINE size), ids offset by rank.  At N > 1 the same line also carries a `strong` record:
the BASELINE corpus (1M u8 rows / 10M binary rows) split over the N ranks, which is what BASELINE.json's metric and
config 3 literally describe; it is latency-bound at 1M / 8 and says so.

`value` is timed with CUDA events on device-resident inputs; `e2e` goes through the host-buffer C-ABI call
(qz_search_topk: H2D query, encode_query, scan + top-k, D2H results) inside the timed region; the roofline entry times
the scan kernel alone with in-stream events (qz_ctx_set_timing).  The default single-GPU run also measures the other
BASELINE configs (binary C3, PQ C4, batched tcgen05 C5, the reference's own 100k x 1024 bench shape C1) as compact
`other_workloads` entries, each with a parity flag against the oracle.

`--impl reference` times the reference's own CPU scorers (oracle/_ref: quantization/cpp/{avx2,sse}.c compiled
unmodified, driven by the restated score_point + caller-side top-k) on the host cores over the same configuration.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (rows per GPU, dim, bytes per row, description)
    "u8": dict(rows=1_000_000, dim=1536, row_bytes=1540, k=100, dtype="u8",
               name="EncodedVectorsU8 dot scan, 1M x 1536-d, single query, top-100 (BASELINE config 2)"),
    "bin": dict(rows=10_000_000, dim=1536, row_bytes=192, k=100, dtype="u1",
                name="EncodedVectorsBin xor-popcount scan, 10M x 1536-bit, top-100 (BASELINE config 3)"),
    "pq": dict(rows=1_000_000, dim=1536, row_bytes=768, k=100, dtype="f32",
               name="EncodedVectorsPQ LUT scan, 1M x 1536-d, chunk 2 (768 codes/vector), top-100 (BASELINE config 4)"),
    # BASELINE config 5 per GPU: 4096 queries x 1.25M rows x 768-d on the tensor cores (rows = 10M / 8 GPUs)
    "u8batch": dict(rows=1_250_000, dim=768, row_bytes=772, k=100, dtype="u8", nq=4096,
                    name="batched EncodedVectorsU8 dot search, 4096 queries x 1.25M rows per GPU x 768-d, top-100 "
                         "per query, tcgen05 int8 (BASELINE config 5, 10M rows over 8 GPUs)"),
}
METRIC = "scored vectors/s (u8 dot, 1M x 1536)"
KERNEL_NAMES = {"u8": "row_scan_kernel<dot,12,full>", "bin": "row_scan_kernel<xor 4 lanes/row,3,full>", "pq": "pq_scan_kernel"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="default 1000 (20 for u8batch)")
    ap.add_argument("--warmup", type=int, default=0, help="default 20 (3 for u8batch)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="u8", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="rows per GPU (default: the workload's BASELINE size)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="headline record: weak (rows per GPU fixed) or strong (the BASELINE corpus split over the GPUs)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="rows of the CPU baseline (default: the full shard)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the other_workloads entries of the default run")
    ap.add_argument("--no-chain", action="store_true", help="ordinary launches instead of chained scans (qz_ctx_set_scan_chaining)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl", "gather"],
                    help="N > 1: top-k exchange inside the scan launch over peer memory (all ranks get the result), "
                         "NCCL all-gather + merge launch, or the rooted peer-memory gather (rank 0 gets the result)")
    a = ap.parse_args()
    a.steps = a.steps or (20 if a.workload == "u8batch" else 1000)
    a.warmup = max(a.warmup or (3 if a.workload == "u8batch" else 20), 3)
    return a


def ncu_traffic(workload):
    """per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/), or None"""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            d = json.load(open(os.path.join(ROOT, "profiles", name))).get(workload, {})
            if d.get("bytes"):
                return d["bytes"], f"profiles/{name} (dram__bytes_read.sum + dram__bytes_write.sum of one committed " \
                                   "ncu --set full capture of this kernel at this size; a constant, not this run's counter)"
        except Exception:
            pass
    return None, None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return {}, 6650.0, "fallback (B200_PROFILING.md)"


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------- GPU arm
def build_index(qz, torch, ctx, wl, rows, seed, id_base, dev):
    """synthetic random embeddings generated on the device (seeded) and encoded there"""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    dim = wl["dim"]
    if wl is WORKLOADS["bin"]:
        vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
        ix = qz.EncodedVectorsBin.empty(vp, ctx=ctx, id_base=id_base)
        step = 1_000_000
        for r in range(0, rows, step):
            n = min(step, rows - r)
            chunk = torch.rand((n, dim), generator=g, device=dev, dtype=torch.float32) - 0.5  # sign = +-1 w.p. 1/2
            ix.encode_rows_dev(chunk.data_ptr(), r, n)
            del chunk
    else:
        data = torch.rand((rows, dim), generator=g, device=dev, dtype=torch.float32)
        vp = qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False)
        if wl is WORKLOADS["pq"]:
            ix = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, 2, ctx=ctx, id_base=id_base)
        else:
            ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx, id_base=id_base)
        del data
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    return ix


def host_query(wl, dim=None):
    rng = np.random.default_rng(43)
    q = rng.random(dim or wl["dim"], dtype=np.float32)
    return q - 0.5 if wl is WORKLOADS["bin"] else q


def make_oracle(O, wl, ix, storage, n):
    vp = ix.vector_parameters
    if wl is WORKLOADS["bin"]:
        return O.OracleBin(storage, n, vp.dim, 16, 0, False), ("reference" if O.ref.available else "port")
    if wl is WORKLOADS["pq"]:
        # score_point_sse is Rust-side intrinsics: restated, not compiled from the reference
        return O.OraclePQ(storage, n, vp.dim, O.pq_division(vp.dim, 2), ix.centroids, 0, False), "port"
    return O.OracleU8(storage, n, vp.dim, 0, False, ix.alpha, ix.offset, ix.multiplier), ("reference" if O.ref.available else "port")


def time_cpu(run, budget_s, min_reps=2):
    run()  # warm (page in the storage)
    reps, t0 = 0, time.perf_counter()
    while True:
        run()
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s and reps >= min_reps:
            return reps, dt


def cpu_baseline(wl, ix, enc_query, gpu_scores, gpu_ids, sample_rows, k, budget_s=6.0):
    """The reference's CPU scorers over rank 0's encoded rows (all of them unless --cpu-sample), timed twice: ONE
    thread — how the reference scans (demos/benches/encode.rs:42-49 is a plain loop) — and all host cores
    (contiguous id ranges, per-thread top-k merged); plus the parity check of the GPU top-k against that scan."""
    import oracle as O
    n = min(sample_rows or ix.count, ix.count)
    rb = ix.row_bytes
    storage = np.ascontiguousarray(ix.download()[: n * rb])
    cores = os.cpu_count() or 1
    o, kind = make_oracle(O, wl, ix, storage, n)
    q = np.ascontiguousarray(enc_query).view(np.float32) if wl is WORKLOADS["pq"] else enc_query
    os_, oi = o.search(q, k, True, threads=cores)
    if n == ix.count and gpu_ids is not None:
        gs, gi = gpu_scores, gpu_ids
        where = "the timed GPU result"
    else:  # a prefix of the shard: re-scan that prefix on the GPU through a sub-index
        import quantization_b200 as qz
        vp = ix.vector_parameters
        svp = qz.VectorParameters(vp.dim, n, vp.distance_type, vp.invert)
        if wl is WORKLOADS["bin"]:
            sub = qz.EncodedVectorsBin(ix.ctx, storage, svp)
        elif wl is WORKLOADS["pq"]:
            sub = qz.EncodedVectorsPQ(ix.ctx, storage, svp, 2, ix.centroids)
        else:
            sub = qz.EncodedVectorsU8(ix.ctx, storage, svp, ix.alpha, ix.offset, ix.multiplier)
        gs, gi = sub.search_top_k(enc_query, k, True)
        sub.close()
        where = "a GPU scan of the same prefix"
    if wl is WORKLOADS["pq"]:
        parity = bool(np.allclose(np.sort(gs), np.sort(os_), rtol=1e-5, atol=0))
    else:
        parity = bool(np.array_equal(gi, oi) and np.array_equal(np.asarray(gs).view(np.uint32), os_.view(np.uint32)))
    r_all, t_all = time_cpu(lambda: o.search(q, k, True, threads=cores), budget_s * 0.4)
    r_one, t_one = time_cpu(lambda: o.search(q, k, True, threads=1), budget_s * 0.6, min_reps=1)
    return {"value": n * r_all / t_all, "unit": "vectors/s", "cores": cores, "kind": kind,
            "value_all_cores": n * r_all / t_all, "value_1_thread": n * r_one / t_one, "nproc": cores,
            "cpu_model": cpu_model(),
            "sample": f"{'all' if n == ix.count else 'first'} {n} rows of rank 0's shard; {r_all} scans + top-{k} on {cores} "
                      f"threads (contiguous id ranges, per-thread top-k merged) in {t_all:.1f} s, {r_one} on 1 thread "
                      f"(how the reference scans: demos/benches/encode.rs:42-49) in {t_one:.1f} s",
            "gpu_topk_matches_cpu_on_sample": parity, "parity_checked_against": where}


class Timed:
    """W warm-up steps, then K steps between two CUDA events on `stream`, a barrier + synchronize on both sides"""

    def __init__(self, torch, dist, stream, world, ctx=None):
        self.torch, self.dist, self.stream, self.world, self.ctx = torch, dist, stream, world, ctx
        self.launches = 0  # kernels of libqz_cuda launched inside the last timed region

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def run(self, step, steps, warmup):
        t = self.torch
        for _ in range(warmup):
            step()
        self.barrier()
        e0, e1 = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
        self.barrier()
        l0 = self.ctx.launch_count if self.ctx is not None else 0
        e0.record(self.stream)
        for _ in range(steps):
            step()
        e1.record(self.stream)
        self.barrier()
        self.launches = (self.ctx.launch_count - l0) if self.ctx is not None else 0
        return e0.elapsed_time(e1)

    def max_over_ranks(self, values, dev):
        tm = self.torch.tensor(values, device=dev, dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(tm, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in tm]


def global_topk_check(torch, dist, ix, d_enc, k, world, rank, dev, peer_scores, peer_ids):
    """N > 1: every rank also runs the plain single-shard scan (qz_scan_topk_dev), the k keys per rank are
    all-gathered over NCCL and rank 0 merges them on the CPU with the oracle's merge (score, then lower id): the
    peer-memory exchange's result must be exactly that (u8.rs:331-348 for the scores themselves)."""
    from quantization_b200.sharded import keys_to_results
    d_keys = torch.zeros(k, dtype=torch.int64, device=dev)
    ix.search_top_k_dev(d_enc.data_ptr(), k, True, d_keys.data_ptr(), 0, 0)
    d_all = torch.zeros(world * k, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(d_all, d_keys)
    torch.cuda.synchronize()
    if rank != 0:
        return None
    import oracle as O
    keys = d_all.cpu().numpy().view(np.uint64)
    s, i = keys_to_results(keys, True)
    ms, mi = O.merge_topk(s, i, k, True)
    n = ms.size
    return bool(np.array_equal(mi, np.asarray(peer_ids)[:n]) and
                np.array_equal(ms.view(np.uint32), np.asarray(peer_scores, np.float32)[:n].view(np.uint32)))


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import quantization_b200 as qz
    from quantization_b200.sharded import ShardedSearch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.workload]
    rows = args.rows or wl["rows"]
    k = wl["k"]
    stream = torch.cuda.Stream(device=dev)  # a real (non-default) stream shared by torch, NCCL and libqz_cuda
    torch.cuda.set_stream(stream)
    ctx = qz.Context(local, stream=stream.cuda_stream)
    # back-to-back device-resident scans are chained (programmatic dependent launch): nothing else is enqueued on this
    # stream between two steps of the timed loop, which is the promise qz_ctx_set_scan_chaining asks for
    ctx.set_scan_chaining(not args.no_chain)
    T = Timed(torch, dist, stream, world, ctx)
    q_host = host_query(wl)
    mode = args.exchange if world > 1 else "none"
    use_peer = mode in ("peer", "gather")

    def measure(ix, rows_here, steps, warmup, with_e2e):
        """device-timed steps, kernel-only steps, optional host-facing e2e steps of one sharded index"""
        enc = ix.encode_query(q_host)                      # host blob through the C-ABI
        d_enc = torch.from_numpy(enc.copy()).to(dev)       # device-resident copy for the kernel-timed loop
        d_s = torch.zeros(k, dtype=torch.float32, device=dev)
        d_i = torch.zeros(k, dtype=torch.int32, device=dev)
        ss = ShardedSearch(ix, device=dev, peer=use_peer, root=(0 if mode == "gather" else None)) if world > 1 else None
        out = {"enc": enc, "d_enc": d_enc}

        def step_dev():
            if world == 1:
                ix.search_top_k_dev(d_enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr())
                return d_s, d_i
            return ss.search_top_k_dev(d_enc.data_ptr(), k, True)

        ms = T.run(step_dev, steps, warmup)
        out["launches"] = T.launches
        # the same loop with ordinary (fully serialised) launches, for the record
        ctx.set_scan_chaining(False)
        plain_steps = max(steps // 4, 20)
        out["plain_ms_per_step"] = T.max_over_ranks([T.run(step_dev, plain_steps, 3) / plain_steps], dev)[0]
        ctx.set_scan_chaining(not args.no_chain)
        # the dominant kernel alone: the same loop again with in-stream events around each scan launch.  (A second,
        # event-bracketed pass: its per-launch mean can differ from ms_per_step by a percent either way.)
        ctx.set_timing(True)
        for _ in range(steps):
            step_dev()
        T.barrier()
        kern_ms, kern_n = ctx.get_timing()
        ctx.set_timing(False)
        rs, ri = step_dev()
        T.barrier()
        out["scores"], out["ids"] = rs.cpu().numpy().copy(), ri.cpu().numpy().astype(np.uint32)
        out["step_dev"] = step_dev
        e2e_ms = 0.0
        if with_e2e:
            # e2e: the host-buffer call a user of the trait makes, every step: f32 query -> encode_query -> scan + top-k
            # (-> exchange) -> (scores, ids) on the host
            call = (lambda: ss.search(q_host, k, True)) if use_peer else (lambda: ix.search(q_host, k, True))
            for _ in range(3):
                call()
            T.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                hs, hi = call()
                if world > 1 and not use_peer:  # host-side exchange of the k results per rank
                    mine = torch.from_numpy(np.concatenate([hs.view(np.uint32), hi]).astype(np.int64)).to(dev)
                    allr = torch.empty(world * mine.numel(), dtype=torch.int64, device=dev)
                    dist.all_gather_into_tensor(allr, mine)
                    allr.cpu()
            T.barrier()
            e2e_ms = (time.perf_counter() - t0) * 1e3
        out["ms"], out["e2e_ms"] = T.max_over_ranks([ms, e2e_ms], dev)
        out["kern_ms"], out["kern_n"] = kern_ms, kern_n
        if world > 1:  # the global result against a CPU merge of the per-shard results
            out["global_ok"] = global_topk_check(torch, dist, ix, d_enc, k, world, rank, dev, out["scores"], out["ids"])
        out["ss"] = ss
        return out

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ix = build_index(qz, torch, ctx, wl, rows, 42 + rank, rank * rows, dev)
    if rank == 0:
        sampler.lines.clear()
    weak = measure(ix, rows, args.steps, args.warmup, True)
    # nvidia-smi samples every 100 ms: if the timed region was shorter, keep the same load running (untimed, the same
    # number of steps on every rank: the peer exchange is collective) until ~0.7 s of samples exist
    extra = int(min(20000, max(0, 700.0 / max(weak["ms"] / args.steps, 1e-3) - 2 * args.steps)))
    for _ in range(extra):
        weak["step_dev"]()
    T.barrier()
    clocks = sampler.stop() if rank == 0 else None

    # single-GPU reference time of the SAME 1M-row (10M-row) shard without any exchange, and the strong-scaling record
    strong = None
    if world > 1 and args.workload in ("u8", "bin") and not args.rows:
        d_s1 = torch.zeros(k, dtype=torch.float32, device=dev)
        d_i1 = torch.zeros(k, dtype=torch.int32, device=dev)
        single_ms = T.run(lambda: ix.search_top_k_dev(weak["d_enc"].data_ptr(), k, True, 0, d_s1.data_ptr(), d_i1.data_ptr()),
                          args.steps, args.warmup)
        (single_ms,) = T.max_over_ranks([single_ms], dev)
        lo, hi = rank * rows // world, (rank + 1) * rows // world
        six = build_index(qz, torch, ctx, wl, hi - lo, 4242 + rank, lo, dev)
        sm = measure(six, hi - lo, args.steps, args.warmup, True)
        t1, tn = single_ms / args.steps, sm["ms"] / args.steps
        strong = {"value": rows * args.steps / (sm["ms"] * 1e-3), "unit": "vectors/s", "ms_per_step": tn,
                  "total_rows": rows, "rows_per_gpu": hi - lo, "single_gpu_ms_per_step_same_run": t1,
                  "efficiency_vs_n1": t1 / (world * tn),
                  "e2e": {"value": rows * args.steps / (sm["e2e_ms"] * 1e-3), "ms_per_step": sm["e2e_ms"] / args.steps},
                  "kernel_ms": sm["kern_ms"] / max(sm["kern_n"], 1),
                  "ms_per_step_unchained": sm["plain_ms_per_step"],
                  "global_topk_matches_merge_of_shards": sm.get("global_ok"),
                  "note": f"the BASELINE corpus split over {world} GPUs: {(hi - lo) * wl['row_bytes'] / 1e6:.0f} MB per GPU "
                          f"= {(hi - lo) * wl['row_bytes'] / 6.5e6:.0f} us of HBM time per step against a fixed cost "
                          "(launch, per-CTA top-k, cross-CTA merge, exchange) of the same order: latency-bound"}
        if sm["ss"] is not None:
            sm["ss"].close()
        six.close()

    if rank == 0:
        pk, peak, peak_src = peaks()
        total_rows = rows * world
        ms, e2e_ms = weak["ms"], weak["e2e_ms"]
        scan_ms = weak["kern_ms"] / max(weak["kern_n"], 1)
        achieved = rows * wl["row_bytes"] / (scan_ms * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(args.workload) if rows == wl["rows"] else (None, None)
        head = {"value": total_rows * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps, "scaling": "weak"}
        if args.scaling == "strong" and strong:
            head = {"value": strong["value"], "ms_per_step": strong["ms_per_step"], "scaling": "strong"}
        out = {
            "metric": METRIC if args.workload == "u8" else f"scored vectors/s ({args.workload} scan)",
            "value": head["value"], "unit": "vectors/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"],
            "higher_is_better": True, "scaling": head["scaling"], "vs_baseline": None, "dtype": wl["dtype"],
            "data": "synthetic: torch.rand embeddings generated and encoded on the device, seed 42 + rank",
            "config": config_of(args, wl, rows, world),
            "notes": {"exchange": {"none": "none (1 GPU)",
                                   "peer": "peer memory inside the scan launch (qz_scan_topk_peer_dev): every rank ends with the global top-k",
                                   "gather": "peer memory inside the scan launch, rooted (qz_scan_topk_gather_dev): rank 0 ends with the global top-k",
                                   "nccl": "NCCL all-gather of k keys + merge launch"}[mode],
                      "launches": ("chained: consecutive scans are launched as programmatic dependents, so the final "
                                   "cross-CTA merge (and exchange) of step i overlaps the streaming of step i+1; every step "
                                   "still scans the whole shard and produces its own top-k" if not args.no_chain else
                                   "ordinary stream-ordered launches (--no-chain)"),
                      "ms_per_step_unchained": weak["plain_ms_per_step"],
                      "l2": f"input {rows * wl['row_bytes'] / 1e6:.0f} MB per GPU per step "
                            + ("exceeds" if rows * wl["row_bytes"] > 126e6 else "FITS IN") + " the 126 MB L2; no flush"},
            "e2e": {"value": total_rows * args.steps / (e2e_ms * 1e-3), "unit": "vectors/s",
                    "h2d_bytes_per_step": int(q_host.nbytes),
                    "d2h_bytes_per_step": int(2 * k * 4),  # one copy: k scores and k ids are contiguous on the device
                    "ms_per_step": e2e_ms / args.steps,
                    "call": ("qz_search_topk_peer" if use_peer else "qz_search_topk") +
                            " (host f32 query -> host scores, ids), one call per step; the query is read from, and the "
                            "results are written to, pinned host memory by the kernels themselves (no copy operations)"},
            "gpu_launches": int(weak["launches"]),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "kernel": KERNEL_NAMES[args.workload],
                         "kernel_ms": scan_ms, "kernel_launches_timed": weak["kern_n"],
                         "kernel_ms_note": "mean of in-stream event pairs around each scan launch in a second pass of the "
                                           "same loop (not the pass ms_per_step comes from)",
                         "peak_source": peak_src, "frac_of_nominal_8TBs": achieved / 8000.0,
                         # the same ratio for the timed loop itself (chained launches: the merge of query i overlaps the
                         # streaming of query i+1, so this is the rate of a stream of queries, not of one kernel)
                         "stream_gbs": rows * wl["row_bytes"] / (ms / args.steps * 1e-3) / 1e9,
                         "stream_frac": rows * wl["row_bytes"] / (ms / args.steps * 1e-3) / 1e9 / peak,
                         "peak_note": "the peak is the driver-measured COPY bandwidth; a read-only stream can exceed it "
                                      "(profiles/r1_hbm_probe.txt: 7.2-7.5 TB/s of bulk-copy reads)"},
            "clocks": clocks,
            "top1_id": int(weak["ids"][0]),
        }
        if world > 1:
            out["global_topk_matches_merge_of_shards"] = weak.get("global_ok")
            out["weak"] = {"value": total_rows * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
                           "rows_per_gpu": rows}
        if strong:
            out["strong"] = strong
        if not args.no_cpu and world == 1:
            out["cpu_baseline"] = cpu_baseline(wl, ix, weak["enc"], weak["scores"], weak["ids"], args.cpu_sample, k)
    if weak["ss"] is not None:
        weak["ss"].close()
    ix.close()
    if rank == 0:
        if world == 1 and args.workload == "u8" and not args.no_others and not args.rows:
            out["other_workloads"] = other_workloads(qz, torch, ctx, T, dev, peak, pk)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def config_of(args, wl, rows, world):
    """the workload description shared verbatim by both arms (the driver compares them)"""
    return {"workload": wl["name"], "rows_per_gpu": rows, "dim": wl["dim"], "k": wl["k"],
            "bytes_per_vector": wl["row_bytes"], "sharding": f"point-id ranges x{world}"}


# ------------------------------------------------------------------------- the other BASELINE configs, compact
def other_workloads(qz, torch, ctx, T, dev, hbm_peak, pk):
    """C3 (binary), C4 (PQ), C5 (batched tcgen05) and C1 (the reference's criterion shape) on one GPU, ~50 steps
    each, device-timed, each with a parity flag against the oracle — so that the driver's default run records them"""
    import oracle as O
    res = {}
    k = 100
    for name, steps in (("bin", 50), ("pq", 50)):
        wl = WORKLOADS[name]
        try:
            ix = build_index(qz, torch, ctx, wl, wl["rows"], 42, 0, dev)
            q = host_query(wl)
            enc = ix.encode_query(q)
            d_enc = torch.from_numpy(enc.copy()).to(dev)
            d_s = torch.zeros(k, dtype=torch.float32, device=dev)
            d_i = torch.zeros(k, dtype=torch.int32, device=dev)
            step = lambda: ix.search_top_k_dev(d_enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr())  # noqa: E731
            ms = T.run(step, steps, 5)
            launches = T.launches
            ctx.set_timing(True)
            for _ in range(steps):
                step()
            T.barrier()
            kms, kn = ctx.get_timing()
            ctx.set_timing(False)
            gs, gi = d_s.cpu().numpy(), d_i.cpu().numpy().astype(np.uint32)
            n = 200_000  # parity on a prefix the CPU scans in well under a second
            cb = cpu_baseline(wl, ix, enc, None, None, n, k, budget_s=2.0)
            ach = wl["rows"] * wl["row_bytes"] / (kms / max(kn, 1) * 1e-3) / 1e9
            res[name] = {"workload": wl["name"], "value": wl["rows"] * steps / (ms * 1e-3), "unit": "vectors/s",
                         "ms_per_step": ms / steps, "steps": steps, "gpu_launches_per_step": launches / steps,
                         "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "frac": ach / hbm_peak,
                                      "kernel": KERNEL_NAMES[name], "kernel_ms": kms / max(kn, 1)},
                         "parity_vs_oracle_on_first_200k_rows": cb["gpu_topk_matches_cpu_on_sample"],
                         "cpu_1_thread": cb["value_1_thread"], "cpu_all_cores": cb["value_all_cores"], "top1_id": int(gi[0])}
            if name == "pq":
                res[name]["lookups_per_s"] = wl["rows"] * 768 / (kms / max(kn, 1) * 1e-3)
            ix.close()
        except Exception as e:  # noqa: BLE001
            res[name] = {"error": repr(e)}
    try:
        res["u8batch"] = batch_record(qz, torch, ctx, T, dev, pk, steps=5, warmup=3, nqs_cpu=4, cpu_rows=100_000, compact=True)
    except Exception as e:  # noqa: BLE001
        res["u8batch"] = {"error": repr(e)}
    try:
        res["c1_100k_x_1024"] = c1_record(qz, torch, ctx, T, dev, O, 1024)
        res["c1_100k_x_128"] = c1_record(qz, torch, ctx, T, dev, O, 128)
    except Exception as e:  # noqa: BLE001
        res["c1_100k_x_1024"] = {"error": repr(e)}
    return res


def c1_record(qz, torch, ctx, T, dev, O, dim):
    """BASELINE config 1 / the reference's criterion shapes (demos/benches/encode.rs:15-16: 100k x 1024, uniform [0,1);
    BASELINE.json says 128): 'score all' (:42-49) and 'random access' (:91-102), device-timed, next to the reference's
    own AVX2 scorer on one host thread over the same encoded bytes"""
    count, k = 100_000, 100
    data = O.uniform(42, count * dim).reshape(count, dim)
    query = O.uniform(43, dim)
    vp = qz.VectorParameters(dim, count, qz.DistanceType.Dot, False)
    g = qz.EncodedVectorsU8.encode(data, qz.VecStorageBuilder(), vp, None, ctx=ctx)
    o = O.OracleU8(np.asarray(g.encoded_vectors), count, dim, 0, False, g.alpha, g.offset, g.multiplier)
    enc = g.encode_query(query)
    d_enc = torch.from_numpy(enc.copy()).to(dev)
    d_all = torch.zeros(count, dtype=torch.float32, device=dev)
    d_s = torch.zeros(k, dtype=torch.float32, device=dev)
    d_i = torch.zeros(k, dtype=torch.int32, device=dev)
    ids = np.random.default_rng(1).integers(0, count, count).astype(np.uint32)
    d_ids = torch.from_numpy(ids.view(np.int32)).to(dev)
    steps = 200
    ms_all = T.run(lambda: g.score_all_dev(d_enc.data_ptr(), d_all.data_ptr()), steps, 5)
    full = d_all.cpu().numpy()
    ms_topk = T.run(lambda: g.search_top_k_dev(d_enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr()), steps, 5)
    ms_rand = T.run(lambda: g.score_points_dev(d_enc.data_ptr(), d_ids.data_ptr(), count, d_all.data_ptr()), steps, 5)
    rnd = d_all.cpu().numpy()
    ref = o.scan(enc)
    parity = bool(np.array_equal(full.view(np.uint32), ref.view(np.uint32)) and np.array_equal(rnd.view(np.uint32), ref[ids].view(np.uint32)))
    os_, oi = o.search(enc, k, True)
    parity = parity and bool(np.array_equal(d_i.cpu().numpy().astype(np.uint32), oi))
    r1, t1 = time_cpu(lambda: o.scan(enc), 1.0)
    g.close()
    return {"workload": f"EncodedVectorsU8 dot, 100k x {dim}-d, single query (the reference's criterion bench shape)",
            "score_all_vectors_per_s": count * steps / (ms_all * 1e-3), "score_all_ms": ms_all / steps,
            "scan_topk_vectors_per_s": count * steps / (ms_topk * 1e-3), "scan_topk_ms": ms_topk / steps,
            "random_access_ids_per_s": count * steps / (ms_rand * 1e-3), "random_access_ms_per_100k_ids": ms_rand / steps,
            "cpu_1_thread_score_all_vectors_per_s": count * r1 / t1, "cpu_kind": "reference" if O.ref.available else "port",
            "parity_vs_oracle_all_rows": parity,
            "note": f"{count * (dim + 4) / 1e6:.0f} MB fits the 126 MB L2 and one launch is a few microseconds: latency-bound"}


# ------------------------------------------------------------------------------------------ batched arm
def int8_peak_probe(torch, dev):
    """a measured dense int8 tensor peak on this GPU: torch._int_mm 8192^3 (cuBLASLt), best of 10.  A probe for the
    roofline denominator only — nothing on the product path calls it."""
    try:
        a = torch.randint(-100, 100, (8192, 8192), dtype=torch.int8, device=dev)
        b = torch.randint(-100, 100, (8192, 8192), dtype=torch.int8, device=dev)
        for _ in range(3):
            torch._int_mm(a, b)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch._int_mm(a, b)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        del a, b
        return 2.0 * 8192 ** 3 / (best * 1e-3) / 1e12
    except Exception:  # noqa: BLE001
        return None


def batch_record(qz, torch, ctx, T, dev, pk, steps, warmup, nqs_cpu, cpu_rows, compact, world=1, rank=0, rows=None,
                 no_cpu=False):
    """BASELINE config 5: a step = one batched search (encode nq queries, int8 contraction on the tensor cores,
    per-query top-k) over the rank's shard, + for N > 1 the all-gather of the [nq][k] results and the
    per-query merge.  value = nq * rows * N / t (scored (query, vector) pairs per second)."""
    import ctypes as C
    import torch.distributed as dist
    wl = WORKLOADS["u8batch"]
    rows, dim, k, nq = rows or wl["rows"], wl["dim"], wl["k"], wl["nq"]
    stream = T.stream
    g = torch.Generator(device=dev)
    g.manual_seed(42 + rank)
    data = torch.rand((rows, dim), generator=g, device=dev, dtype=torch.float32)
    ix = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), qz.VectorParameters(dim, rows, qz.DistanceType.Dot, False), None,
                                        ctx=ctx, id_base=rank * rows)
    del data
    q_host = np.random.default_rng(43).random((nq, dim), dtype=np.float32)
    q_pin = torch.from_numpy(q_host).pin_memory()
    d_q = q_pin.to(dev)
    d_s = torch.zeros((nq, k), dtype=torch.float32, device=dev)
    d_i = torch.zeros((nq, k), dtype=torch.int32, device=dev)
    d_as = torch.zeros((world, nq, k), dtype=torch.float32, device=dev)
    d_ai = torch.zeros((world, nq, k), dtype=torch.int32, device=dev)
    d_ms = torch.zeros((nq, k), dtype=torch.float32, device=dev)
    d_mi = torch.zeros((nq, k), dtype=torch.int32, device=dev)
    L = ctx._L

    def step_dev(dq):
        ix.search_batch_dev(dq.data_ptr(), nq, k, True, d_s.data_ptr(), d_i.data_ptr())
        if world > 1:
            dist.all_gather_into_tensor(d_as, d_s)
            dist.all_gather_into_tensor(d_ai, d_i)
            ctx.check(L.qz_merge_topk_batch_dev(ctx.handle, C.c_void_p(d_as.data_ptr()), C.c_void_p(d_ai.data_ptr()), world,
                                                nq, k, 1, C.c_void_p(d_ms.data_ptr()), C.c_void_p(d_mi.data_ptr())))

    sampler = ClockSampler(dev.index or 0)
    if rank == 0 and not compact:
        sampler.start()
    ms = T.run(lambda: step_dev(d_q), steps, warmup)
    launches = T.launches
    ctx.set_timing(True)
    for _ in range(steps):
        step_dev(d_q)
    T.barrier()
    kern_ms, kern_n = ctx.get_timing()
    ctx.set_timing(False)
    clocks = sampler.stop() if rank == 0 and not compact else None
    # e2e: pinned host queries -> H2D -> search -> D2H of the [nq][k] results, every step
    h_s = torch.empty((nq, k), dtype=torch.float32).pin_memory()
    h_i = torch.empty((nq, k), dtype=torch.int32).pin_memory()
    d_q2 = torch.empty_like(d_q)
    T.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        d_q2.copy_(q_pin, non_blocking=True)
        step_dev(d_q2)
        h_s.copy_(d_ms if world > 1 else d_s, non_blocking=True)
        h_i.copy_(d_mi if world > 1 else d_i, non_blocking=True)
        torch.cuda.synchronize()
    T.barrier()
    e2e_s = time.perf_counter() - t0
    ms, e2e_ms = T.max_over_ranks([ms, e2e_s * 1e3], dev)
    if rank != 0:
        ix.close()
        return None
    bf16 = float(pk.get("bf16_tflops", 1590.0))
    int8_meas = int8_peak_probe(torch, dev)
    peak = int8_meas or 2.0 * bf16
    ad = ix.actual_dim
    ops = 2.0 * nq * rows * ad
    gemm_ms = kern_ms / max(steps, 1)  # all contraction launches of one step
    achieved = ops / (gemm_ms * 1e-3) / 1e12
    pairs = float(nq) * rows * world
    out = {
        "metric": "scored (query, vector) pairs/s (batched u8 dot, 4096 q x 10M x 768-d)",
        "value": pairs * steps / (ms * 1e-3), "unit": "vectors/s", "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic: torch.rand embeddings generated and encoded on the "
        "device, seed 42 + rank; numpy default_rng(43) queries",
        "config": {"workload": wl["name"], "rows_per_gpu": rows, "dim": dim, "k": k, "queries": nq,
                   "sharding": f"point-id ranges x{world}"},
        "notes": {"l2": f"codes {rows * ad / 1e6:.0f} MB per GPU exceed the 126 MB L2; no flush"},
        "e2e": {"value": pairs * steps / (e2e_ms * 1e-3), "unit": "vectors/s",
                "h2d_bytes_per_step": int(q_host.nbytes), "d2h_bytes_per_step": int(nq * k * 8),
                "ms_per_step": e2e_ms / steps},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                     "traffic": None, "kernel": "u8_batch_kernel (tcgen05.mma kind::i8), all launches of a step",
                     "kernel_ms": gemm_ms, "kernel_launches_timed": kern_n,
                     "peak_source": ("measured on this GPU: torch._int_mm (cuBLASLt) int8 8192^3, best of 10" if int8_meas else
                                     "2 x measured bf16 burst (MEASURED_PEAKS.json bf16_tflops): the int8 probe did not run"),
                     "int8_ops": "2*Q*N*K", "peak_int8_measured": int8_meas, "peak_2x_bf16_measured": 2.0 * bf16,
                     "frac_of_2x_bf16_measured": achieved / (2.0 * bf16), "frac_of_nominal_4500": achieved / 4500.0},
        "clocks": clocks,
    }
    if not no_cpu:
        import oracle as O
        n = min(cpu_rows, ix.count)
        storage = np.ascontiguousarray(ix.download()[: n * ix.row_bytes])
        o = O.OracleU8(storage, n, dim, 0, False, ix.alpha, ix.offset, ix.multiplier)
        cores = os.cpu_count() or 1
        encs = [ix.encode_query(q_host[i]) for i in range(nqs_cpu)]
        sub = qz.EncodedVectorsU8(ctx, storage, qz.VectorParameters(dim, n, qz.DistanceType.Dot, False), ix.alpha,
                                  ix.offset, ix.multiplier)
        gs, gi = sub.search_batch(q_host[:nqs_cpu], k, True)
        par = True
        t0 = time.perf_counter()
        reps = 0
        while True:
            for i in range(nqs_cpu):
                os_, oi = o.search(encs[i], k, True, threads=cores)
                if reps == 0:
                    par = par and bool(np.array_equal(gi[i], oi) and np.array_equal(gs[i].view(np.uint32), os_.view(np.uint32)))
            reps += 1
            dt = time.perf_counter() - t0
            if dt >= (1.0 if compact else 3.0):
                break
        t1 = time.perf_counter()
        o.search(encs[0], k, True, threads=1)
        one = n / (time.perf_counter() - t1)
        sub.close()
        out["cpu_baseline"] = {"value": n * nqs_cpu * reps / dt, "unit": "vectors/s", "cores": cores,
                               "value_all_cores": n * nqs_cpu * reps / dt, "value_1_thread": one, "nproc": cores,
                               "cpu_model": cpu_model(),
                               "kind": "reference" if O.ref.available else "port",
                               "sample": f"first {n} rows of rank 0's shard x {nqs_cpu} queries, {reps} rounds of "
                                         f"single-query scans + top-{k} (the reference has no batched path), {cores} threads, {dt:.1f} s wall",
                               "gpu_topk_matches_cpu_on_sample": par}
    ix.close()
    if compact:
        return {"workload": wl["name"], "value": out["value"], "unit": "scored pairs/s", "ms_per_step": out["ms_per_step"],
                "steps": steps, "gpu_launches_per_step": launches / steps,
                "roofline": {key: out["roofline"][key] for key in ("bound", "achieved", "peak", "frac", "kernel_ms", "peak_source",
                                                                    "peak_int8_measured", "frac_of_2x_bf16_measured",
                                                                    "frac_of_nominal_4500")},
                "e2e_ms_per_step": out["e2e"]["ms_per_step"],
                "parity_vs_oracle": out.get("cpu_baseline", {}).get("gpu_topk_matches_cpu_on_sample")}
    return out


def run_gpu_batch(args):
    import torch
    import torch.distributed as dist
    import quantization_b200 as qz

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = qz.Context(local, stream=stream.cuda_stream)
    T = Timed(torch, dist, stream, world, ctx)
    pk, _, _ = peaks()
    out = batch_record(qz, torch, ctx, T, dev, pk, args.steps, args.warmup, 8, args.cpu_sample or 200_000, False, world, rank,
                       args.rows or None, args.no_cpu or world > 1)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------- reference arm
def reference_index(O, workload, n, dim):
    """the workload's corpus, encoded by the oracle (the restated reference encode), built in 100k-row pieces so that
    the f32 source (6 GB at 1M x 1536) never exists at once.  u8 of uniform [0, 1): min / max are taken over the
    whole corpus first, as EncodedVectorsU8::encode does (u8.rs:57)."""
    import ctypes as C
    from oracle import oracle as OM
    piece = 100_000
    if workload in ("u8", "u8batch"):
        mn, mx = np.float32(np.finfo(np.float32).max), np.float32(np.finfo(np.float32).min)
        for r in range(0, n, piece):
            m = min(piece, n - r)
            a, b = O.minmax(O.uniform(42, m * dim, first=r * dim))
            mn, mx = min(mn, np.float32(a)), max(mx, np.float32(b))
        alpha, offset = C.c_float(), C.c_float()
        OM._lib.qzo_u8_alpha_offset(C.c_float(mn), C.c_float(mx), C.byref(alpha), C.byref(offset))
        ad = O.u8_actual_dim(dim)
        storage = np.empty(n * (ad + 4), np.uint8)
        for r in range(0, n, piece):
            m = min(piece, n - r)
            rows = O.uniform(42, m * dim, first=r * dim)
            out = storage[r * (ad + 4):(r + m) * (ad + 4)]
            OM._lib.qzo_u8_encode(OM._p(rows, OM._f32p), C.c_uint64(m), C.c_uint32(dim), C.c_int(0), C.c_int(0), alpha, offset,
                                  OM._p(out, OM._u8p))
        mult = OM._lib.qzo_u8_multiplier(alpha, C.c_int(0), C.c_int(0))
        o = O.OracleU8(storage, n, dim, 0, False, alpha.value, offset.value, mult)
        return o, o.encode_query(O.uniform(43, dim)), ("reference" if O.ref.available else "port")
    if workload == "bin":
        rb = O.bin_row_bytes(dim, 16)
        storage = np.empty(n * rb, np.uint8)
        for r in range(0, n, piece):
            m = min(piece, n - r)
            part = O.OracleBin.encode(O.uniform(42, m * dim, -0.5, 0.5, first=r * dim).reshape(m, dim), 0, False, 16)
            storage[r * rb:(r + m) * rb] = part.storage
        o = O.OracleBin(storage, n, dim, 16, 0, False)
        return o, o.encode_query(O.uniform(43, dim, -0.5, 0.5)), ("reference" if O.ref.available else "port")
    bounds = O.pq_division(dim, 2)
    cen = O.uniform(7, 256 * dim).reshape(256, dim)  # scoring cost does not depend on centroid quality
    nch = bounds.size - 1
    codes = np.empty(n * nch, np.uint8)
    for r in range(0, n, piece):
        m = min(piece, n - r)
        codes[r * nch:(r + m) * nch] = O.OraclePQ.assign(O.uniform(42, m * dim, first=r * dim).reshape(m, dim), bounds, cen)
    o = O.OraclePQ(codes, n, dim, bounds, cen, 0, False)
    return o, o.encode_query(O.uniform(43, dim)), "port"


def run_reference(args):
    """The reference's own CPU implementation of the path on the host cores: its C scorers compiled unmodified
    (oracle/_ref) driven by the restated score_point + caller-side top-k, all host threads.  A step = one scan +
    top-k of ONE shard of the configured size (1M rows for the headline workload: the same configuration as the GPU
    arm; the binary workload's 10M-row shard is sampled at 2M rows to keep the arm within minutes — the rate does not
    depend on the row count once the working set exceeds the caches)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import oracle as O
    wl = WORKLOADS[args.workload]
    full = args.rows or wl["rows"]
    n = args.cpu_sample or (min(full, 2_000_000) if args.workload == "bin" else (200_000 if args.workload == "u8batch" else full))
    dim, k = wl["dim"], wl["k"]
    cores = os.cpu_count() or 1
    o, q, kind = reference_index(O, args.workload, n, dim)
    run = lambda: o.search(q, k, True, threads=cores)  # noqa: E731
    for _ in range(args.warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run()
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    t1 = time.perf_counter()
    o.search(q, k, True, threads=1)
    one = n / (time.perf_counter() - t1)
    sample = (f"{n} x {dim} rows ({'the full shard' if n == full else 'a sample of the ' + str(full) + '-row shard'}, encoded by the "
              f"oracle), {args.steps} scans + top-{k}, {cores} threads")
    print(json.dumps({
        "impl": "reference",
        "metric": METRIC if args.workload == "u8" else f"scored vectors/s ({args.workload} scan)",
        "value": v, "unit": "vectors/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": wl["dtype"], "data": "synthetic: SplitMix64 uniform embeddings, seed 42",
        "config": config_of(args, wl, full, args.gpus),
        "cpu_baseline": {"value": v, "unit": "vectors/s", "cores": cores, "kind": kind, "sample": sample,
                         "value_all_cores": v, "value_1_thread": one, "nproc": cores, "cpu_model": cpu_model()},
        "e2e": {"value": v, "unit": "vectors/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "u8batch":
        run_gpu_batch(a)
    else:
        run_gpu(a)
