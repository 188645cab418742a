"""bsz-1 FFN value product (SURVEY a-6): achieved GB/s over the bytes of the rows actually touched, ours vs the reference's own
kernel (oracle/_ref), at the 7.2B shape (D = 16384, C = 4096) for several activation densities; plus the whole bsz-1 step
with the sparse product vs the dense one (B200_NO_SPMV=1).  Usage (GPU box): python tools/bench_spmv.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200 import ops


def timed(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
        with torch.cuda.graph(g, stream=s):
            for _ in range(n):
                fn()
    torch.cuda.current_stream().wait_stream(s)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / n


def main():
    D, C, nmat = 16384, 4096, 24            # 24 different matrices (3.2 GB) so that nothing is served from L2
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
    mats = [(torch.randn(D, C, device="cuda") * 0.05).half() for _ in range(nmat)]
    out = torch.empty(C, dtype=torch.float16, device="cuda")
    res = []
    for density in (0.05, 0.15, 0.5, 1.0):
        vec = torch.randn(D, device="cuda").abs()
        vec[torch.rand(D, device="cuda") > density] = 0
        vec = vec.half()
        nnz = int((vec != 0).sum())
        state = {"i": 0}

        def ours():
            state["i"] = (state["i"] + 1) % nmat
            ops.spmv_forward(vec, mats[state["i"]], out=out)

        t = timed(ours)
        rec = {"density": density, "nnz_rows": nnz, "touched_MB": nnz * C * 2 / 1e6, "ours_us": t * 1e6,
               "ours_GBps_of_touched_rows": nnz * C * 2 / t / 1e9, "frac_of_measured_hbm_peak": nnz * C * 2 / t / 1e9 / peak}
        try:
            from oracle import rwkv7_oracle as O
            lib = O.ref_cuda_lib()
            z = torch.zeros(C, dtype=torch.float16, device="cuda")
            import time
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for i in range(200):
                lib.ref_spmv(D, C, vec.data_ptr(), mats[i % nmat].data_ptr(), z.data_ptr())
            torch.cuda.synchronize()
            tr = (time.perf_counter() - t0) / 200
            rec.update({"reference_kernel_us": tr * 1e6, "reference_GBps_of_touched_rows": nnz * C * 2 / tr / 1e9})
        except Exception as e:  # noqa: BLE001
            rec["reference_kernel"] = repr(e)[:100]
        print(json.dumps(rec), flush=True)
        res.append(rec)


if __name__ == "__main__":
    main()
