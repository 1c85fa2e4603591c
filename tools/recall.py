"""Recall harness in the shape of the reference's test_knn (demos/src/ann_benchmark_data.rs:123-185): per query
encode_query + a scan of every point + the 30 best, then same_10 / same_20 / same_30 against the exact neighbours,
and the timing lines of print_timings (:203-220: min / avg / p95 / p99 / max).

    python tools/recall.py                                  # synthetic: 1M x 256 Gaussian mixture, 200 queries
    python tools/recall.py --rows 100000 --dim 128 --queries 50
    python tools/recall.py --train base.fvecs --test query.fvecs --neighbors gt.ivecs --metric l2
    python tools/recall.py --train train.npy --test test.npy [--neighbors neighbors.npy] --metric cosine

On-disk input is the train / test / neighbors triplet of the ann-benchmarks files the reference downloads
(ann_benchmark_data.rs:23-60), as .npy arrays or TexMex .fvecs / .ivecs (the HDF5 container itself is not read: h5py is
not in the image and the files need the network).  Without --neighbors the exact neighbours are computed with an f32
matmul on the device (torch; harness only, not the product path).  --metric cosine normalises rows and queries
(the reference's cosine_preprocess, :223-233) and scores with Dot; l2 uses DistanceType::L2 and smallest-first."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def read_vecs(path):
    """.npy, or TexMex .fvecs / .ivecs / .bvecs: every record is [int32 d][d values]"""
    if path.endswith(".npy"):
        return np.load(path)
    dt = {"fvecs": np.float32, "ivecs": np.int32, "bvecs": np.uint8}[path.rsplit(".", 1)[-1]]
    raw = np.fromfile(path, np.uint8)
    d = int(raw[:4].view(np.int32)[0])
    rec = 4 + d * np.dtype(dt).itemsize
    if raw.size % rec:
        raise ValueError(f"{path}: size {raw.size} is not a multiple of the record size {rec}")
    return np.ascontiguousarray(raw.reshape(-1, rec)[:, 4:]).view(dt).reshape(-1, d)


def write_vecs(path, a):
    """inverse of read_vecs (used by the test to make a fixture)"""
    if path.endswith(".npy"):
        return np.save(path, a)
    a = np.ascontiguousarray(a)
    head = np.full((a.shape[0], 1), a.shape[1], np.int32).view(np.uint8)
    np.concatenate([head, a.view(np.uint8).reshape(a.shape[0], -1)], axis=1).tofile(path)


def synthetic(torch, dev, rows, dim, nq, seed=42):
    """a Gaussian mixture on the unit sphere (cosine = dot after normalisation)"""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    centers = torch.randn((1000, dim), generator=g, device=dev)
    data = centers[torch.randint(0, 1000, (rows,), generator=g, device=dev)] + 0.6 * torch.randn((rows, dim), generator=g, device=dev)
    queries = centers[torch.randint(0, 1000, (nq,), generator=g, device=dev)] + 0.6 * torch.randn((nq, dim), generator=g, device=dev)
    return data / data.norm(dim=1, keepdim=True), queries / queries.norm(dim=1, keepdim=True)


def exact_neighbors(torch, data, queries, metric, kk=30):
    kk = min(kk, data.shape[0])
    if metric == "l2":
        d = (queries * queries).sum(1, keepdim=True) - 2 * queries @ data.T + (data * data).sum(1)[None, :]
        return torch.topk(d, kk, dim=1, largest=False).indices.cpu().numpy()
    return torch.topk(queries @ data.T, kk, dim=1).indices.cpu().numpy()


def print_timings(t, out=print):
    """ann_benchmark_data.rs:203-220"""
    t = np.sort(np.asarray(t, np.float64))
    if not t.size:
        return {}
    r = {"min": t[0], "avg": t.mean(), "p95": t[int(np.float32(t.size) * np.float32(0.95))],
         "p99": t[int(np.float32(t.size) * np.float32(0.99))], "max": t[-1]}
    out(f"  Min search time: {r['min']:.3f}ms\n  Avg search time: {r['avg']:.3f}ms\n  p95 search time: {r['p95']:.3f}ms\n"
        f"  p99 search time: {r['p99']:.3f}ms\n  Max search time: {r['max']:.3f}ms")
    return r


def test_knn(name, enc, q_host, neighbors, largest, out=print):
    """ann_benchmark_data.rs:123-185 with the scan + heap replaced by one search call"""
    same = np.zeros(3)
    timings = []
    for j in range(q_host.shape[0]):
        t0 = time.perf_counter()
        _, ids = enc.search(q_host[j], 30, largest)  # encode_query + scan + top-30
        timings.append((time.perf_counter() - t0) * 1e3)
        truth = set(np.asarray(neighbors[j]).tolist())
        for n, kk in enumerate((10, 20, 30)):
            same[n] += len(set(ids[:kk].tolist()) & truth)  # same_count(&knn[0..kk], &neighbors[j])
    same /= q_host.shape[0]
    out(f"{name}\n  queries count: {q_host.shape[0]}\n  same_10: {same[0]:.2f}\n  same_20: {same[1]:.2f}\n  same_30: {same[2]:.2f}")
    return {"name": name, "same_10": same[0], "same_20": same[1], "same_30": same[2], "timings": print_timings(timings, out)}


def run(train=None, test=None, neighbors=None, metric="cosine", rows=1_000_000, dim=256, queries=200, codecs=("u8", "u8q", "bin", "pq"),
        pq_chunk=2, out=print):
    import torch
    import quantization_b200 as qz
    dev = torch.device("cuda", 0)
    if train:
        data = torch.from_numpy(np.ascontiguousarray(read_vecs(train), np.float32)).to(dev)
        q = torch.from_numpy(np.ascontiguousarray(read_vecs(test), np.float32)).to(dev)[:queries]
        if metric == "cosine":
            data = data / data.norm(dim=1, keepdim=True).clamp_min(1e-12)
            q = q / q.norm(dim=1, keepdim=True).clamp_min(1e-12)
        src = f"{train} / {test}"
    else:
        data, q = synthetic(torch, dev, rows, dim, queries)
        src = "synthetic Gaussian mixture on the unit sphere"
    rows, dim = data.shape
    nb = read_vecs(neighbors)[:q.shape[0]] if neighbors else exact_neighbors(torch, data, q, metric)
    out(f"data: {src}: {rows} x {dim}, {q.shape[0]} queries, metric {metric}, neighbours from "
        + (neighbors or "an exact f32 scan"))
    largest = metric != "l2"
    dist = qz.DistanceType.L2 if metric == "l2" else qz.DistanceType.Dot
    vp = qz.VectorParameters(dim, rows, dist, False)
    ctx = qz.Context(0)
    q_host = q.cpu().numpy()
    res = []
    data = data.contiguous()
    for c in codecs:
        t0 = time.perf_counter()
        if c == "u8":
            enc, name = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, None, ctx=ctx), "u8 (no quantile)"
        elif c == "u8q":
            enc, name = qz.EncodedVectorsU8.encode_dev(data.data_ptr(), vp, 0.99, ctx=ctx), "u8 (quantile 0.99)"
        elif c == "bin":
            enc, name = qz.EncodedVectorsBin.empty(vp, ctx=ctx), "binary (1 bit / dimension)"
            enc.encode_rows_dev(data.data_ptr(), 0, rows)
        else:
            enc, name = qz.EncodedVectorsPQ.encode_dev(data.data_ptr(), vp, pq_chunk, ctx=ctx), f"pq (chunk {pq_chunk})"
        torch.cuda.synchronize()
        out(f"encode {name}: {(time.perf_counter() - t0) * 1e3:.1f} ms")
        res.append(test_knn(name, enc, q_host, nb, largest, out))
        enc.close()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--train"); ap.add_argument("--test"); ap.add_argument("--neighbors")
    ap.add_argument("--metric", default="cosine", choices=["cosine", "dot", "l2"])
    ap.add_argument("--rows", type=int, default=1_000_000); ap.add_argument("--dim", type=int, default=256)
    ap.add_argument("--queries", type=int, default=200)
    ap.add_argument("--codecs", default="u8,u8q,bin,pq")
    a = ap.parse_args()
    if bool(a.train) != bool(a.test):
        ap.error("--train and --test go together")
    run(a.train, a.test, a.neighbors, a.metric, a.rows, a.dim, a.queries, tuple(a.codecs.split(",")))
