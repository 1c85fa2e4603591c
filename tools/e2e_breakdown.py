"""Where the host-facing call's time goes beyond the scan kernel (C2 by default): wall-clock microseconds per call,
median of `reps` calls, each followed by a stream synchronisation.

    python tools/e2e_breakdown.py [u8|bin|pq] [rows]

  sync            qz_ctx_synchronize on an idle stream (binding call + driver floor)
  encode+sync     qz_encode_query_dev reading the f32 query from PINNED HOST memory, then synchronise
  scan+sync       qz_scan_topk_dev on a device-resident encoded query, results to device memory, then synchronise
  scan->host+sync the same with scores / ids stored to pinned host memory by the kernel
  search          qz_search_topk through the Python mirror (what bench.py's e2e times)
  search (raw)    the same C call with every ctypes argument prepared once (what a compiled caller pays)
"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench as B
import quantization_b200 as qz


def med_us(f, reps=400, warm=20):
    for _ in range(warm):
        f()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    ts = np.array(ts) * 1e6
    return float(np.median(ts)), float(np.percentile(ts, 10)), float(np.percentile(ts, 90))


def main():
    wl_name = sys.argv[1] if len(sys.argv) > 1 else "u8"
    wl = B.WORKLOADS[wl_name]
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else wl["rows"]
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = qz.Context(0, stream=stream.cuda_stream)
    k = wl["k"]
    ix = B.build_index(qz, torch, ctx, wl, rows, 42, 0, dev)
    q_host = B.host_query(wl)
    q_pin = torch.from_numpy(q_host.copy()).pin_memory()
    enc = ix.encode_query(q_host)
    d_enc = torch.from_numpy(enc.copy()).to(dev)
    d_blob = torch.zeros(max(ix.query_bytes, 16), dtype=torch.uint8, device=dev)
    d_s = torch.zeros(k, dtype=torch.float32, device=dev)
    d_i = torch.zeros(k, dtype=torch.int32, device=dev)
    h_out = torch.zeros(2 * k, dtype=torch.float32).pin_memory()
    sync = ctx.synchronize

    def enc_sync():
        ix.encode_query_dev(q_pin.data_ptr(), 1, d_blob.data_ptr())
        sync()

    def scan_sync():
        ix.search_top_k_dev(d_enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr())
        sync()

    def scan_host_sync():
        ix.search_top_k_dev(d_enc.data_ptr(), k, True, 0, h_out.data_ptr(), h_out.data_ptr() + 4 * k)
        sync()

    L = ctx._L
    scores, ids, n = np.empty(k, np.float32), np.empty(k, np.uint32), C.c_uint32(0)
    args = (ix.handle, C.c_void_p(q_host.ctypes.data), k, 1, C.c_void_p(scores.ctypes.data), C.c_void_p(ids.ctypes.data),
            C.byref(n))

    def search_raw():
        L.qz_search_topk(*args)

    # kernel alone, for reference (event-timed, unchained)
    ctx.set_scan_chaining(False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(20):
        ix.search_top_k_dev(d_enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr())
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(200):
        ix.search_top_k_dev(d_enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
    print(f"workload {wl_name}, {rows} rows, k {k}: scan kernel back to back {e0.elapsed_time(e1) * 1000 / 200:.1f} us per launch")
    for name, f in (("sync", sync), ("encode+sync", enc_sync), ("scan+sync", scan_sync), ("scan->host+sync", scan_host_sync),
                    ("search", lambda: ix.search(q_host, k, True)), ("search (raw)", search_raw)):
        m, lo, hi = med_us(f)
        print(f"  {name:16s} {m:8.1f} us   (p10 {lo:.1f}, p90 {hi:.1f})")
    rs, ri = ix.search(q_host, k, True)
    assert np.array_equal(ri, ids[: n.value]) and np.array_equal(rs.view(np.uint32), scores[: n.value].view(np.uint32))
    assert np.array_equal(d_i.cpu().numpy().view(np.uint32), ri)
    ix.close()


if __name__ == "__main__":
    main()
