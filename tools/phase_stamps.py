"""Per-phase timing of the single-query scan kernel from its in-kernel stamps (qz_ctx_set_phase_stamps).

    python tools/phase_stamps.py [u8|bin] [rows ...]

For each shard size: the event-timed step (launch to launch, 500 steps; plain and with chained launches), and from the stamps of 50 separately
synchronised launches the medians of: kernel span (first CTA in -> last CTA out, globaltimer), per-CTA warm-up,
streaming, local list, and the last CTA's final merge (+ peer exchange when there is one).  Cycle counts are
converted with the SM clock read from NVML during the run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import quantization_b200 as qz
import bench as B

def sm_mhz():
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)
        return float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
    except Exception:
        return 1920.0

def main():
    wl_name = sys.argv[1] if len(sys.argv) > 1 else "u8"
    wl = B.WORKLOADS[wl_name]
    sizes = [int(x) for x in sys.argv[2:]] or ([125_000, 250_000, 500_000, 1_000_000] if wl_name == "u8" else [1_250_000, 2_500_000, 5_000_000, 10_000_000])
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
    ctx = qz.Context(0, stream=stream.cuda_stream)
    k = wl["k"]
    print(f"workload {wl_name}, k {k}, {ctx.sm_count} SMs")
    for rows in sizes:
        ix = B.build_index(qz, torch, ctx, wl, rows, 42, 0, dev)
        enc = torch.from_numpy(np.frombuffer(ix.encode_query(B.host_query(wl)), np.uint8).copy()).to(dev)
        d_s = torch.zeros(k, dtype=torch.float32, device=dev); d_i = torch.zeros(k, dtype=torch.int32, device=dev)
        step = lambda: ix.search_top_k_dev(enc.data_ptr(), k, True, 0, d_s.data_ptr(), d_i.data_ptr())
        def timed(chained):
            ctx.set_scan_chaining(chained)
            for _ in range(20): step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record(stream)
            for _ in range(500): step()
            e1.record(stream); torch.cuda.synchronize()
            return e0.elapsed_time(e1) * 1000 / 500
        step_us, chained_us = timed(False), timed(True)
        ctx.set_scan_chaining(False)
        stamps = torch.zeros((ctx.sm_count, 8), dtype=torch.int64, device=dev)
        ctx.set_phase_stamps(stamps.data_ptr())
        rec = []
        mhz = []
        for _ in range(50):
            step(); torch.cuda.synchronize()
            s = stamps.cpu().numpy().astype(np.int64)
            s = s[s[:, 0] != 0]
            mhz.append(sm_mhz())
            last = s[s[:, 5] != 0]
            rec.append(dict(span=(s[:, 7].max() - s[:, 0].min()) / 1000.0, entry_skew=(s[:, 0].max() - s[:, 0].min()) / 1000.0,
                            warm=np.median(s[:, 2] - s[:, 1]), stream=np.median(s[:, 3] - s[:, 2]), stream_max=(s[:, 3] - s[:, 2]).max(),
                            fin=np.median(s[:, 4] - s[:, 3]), body_max=(s[:, 4] - s[:, 1]).max(),
                            merge=(last[0, 5] - last[0, 4]) if len(last) else 0, xchg=(last[0, 6] - last[0, 5]) if len(last) else 0,
                            ctas=len(s)))
        ctx.set_phase_stamps(0)
        f = np.median(mhz)
        med = lambda key: float(np.median([r[key] for r in rec]))
        cyc = lambda key: med(key) / f
        ideal = rows * wl["row_bytes"] / B.peaks()[1] / 1e3
        print(f"rows {rows:>9}: step {step_us:6.1f} us, chained {chained_us:6.1f} us | ideal (bytes / measured HBM peak) {ideal:6.1f} us | kernel span {med('span'):6.1f} us, CTA entry skew {med('entry_skew'):4.1f} | "
              f"per CTA: warm-up {cyc('warm'):5.1f}, stream {cyc('stream'):6.1f} (max {cyc('stream_max'):6.1f}), local list {cyc('fin'):4.1f}, entry->list max {cyc('body_max'):6.1f} | "
              f"last CTA: merge {cyc('merge'):4.1f}, exchange {cyc('xchg'):4.1f} | CTAs {int(med('ctas'))}, SM {f:.0f} MHz")
        del ix

main()
