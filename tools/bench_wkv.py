"""Micro-benchmark of the WKV-7 decode kernel alone (T = 1): achieved algorithmic GB/s vs the measured HBM peak.
Usage (GPU box): python tools/bench_wkv.py [--geom 7.2B] [--bsz 256] [--ref]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from chirrup_b200 import ops
from chirrup_b200.synth import GEOMETRIES


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="7.2B")
    ap.add_argument("--bsz", type=int, default=256)
    ap.add_argument("--T", type=int, default=1)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ref", action="store_true", help="also time the reference kernel from oracle/_ref")
    a = ap.parse_args()
    g = GEOMETRIES[a.geom]
    B, T, C, H, L = a.bsz, a.T, g.n_embd, g.n_head, g.n_layer
    dev = "cuda"
    # all layers' state so that each launch streams a different 4 GiB/L slab: no L2 reuse between launches
    state = (torch.randn(L, B, H, 64, 64, device=dev) * 0.1).half()
    mk = lambda s=0.5: (torch.randn(B, T, C, device=dev) * s).half()
    r, k, v = mk(), mk(), mk()
    w = (torch.randn(B, T, C, device=dev) - 1).half()
    kk = torch.nn.functional.normalize(torch.randn(B, T, H, 64, device=dev), dim=-1).view(B, T, C)
    aa, bb = (-kk).half(), (kk * torch.sigmoid(torch.randn(B, T, C, device=dev))).half()
    et = torch.zeros(B, dtype=torch.int32, device=dev)
    y = torch.empty_like(r)
    bytes_per_launch = B * (2 * H * 64 * 64 * 2) + B * T * 7 * C * 2 + B * 4

    def run_ours():
        for l in range(L):
            ops.wkv7_forward_seq(state[l], r, w, k, v, aa, bb, et, y=y)

    def run_ref():
        from oracle import rwkv7_oracle as O
        lib = O.ref_cuda_lib()
        for l in range(L):
            lib.ref_seq(B, T, C, H, state[l].data_ptr(), r.data_ptr(), w.data_ptr(), k.data_ptr(), v.data_ptr(),
                        aa.data_ptr(), bb.data_ptr(), y.data_ptr(), et.data_ptr())

    res = {}
    for name, fn in [("ours", run_ours)] + ([("reference_kernel", run_ref)] if a.ref else []):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if name == "reference_kernel":   # launches on the legacy default stream (reference .cu:315) = torch's default stream:
            ds = torch.cuda.default_stream()          # events there bracket them; the launches cannot be graph-captured
            s.record(ds)
            for _ in range(a.iters):
                fn()
            e.record(ds)
            torch.cuda.synchronize()
            ms = s.elapsed_time(e) / a.iters
        else:                              # graph replay: the per-launch Python cost (~15 us) must not be what is timed
            g = torch.cuda.CUDAGraph()
            st = torch.cuda.Stream()
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                fn()
                with torch.cuda.graph(g, stream=st):
                    fn()
            torch.cuda.current_stream().wait_stream(st)
            g.replay()
            torch.cuda.synchronize()
            s.record()
            for _ in range(a.iters):
                g.replay()
            e.record()
            torch.cuda.synchronize()
            ms = s.elapsed_time(e) / a.iters
        gbs = bytes_per_launch * L / ms / 1e6
        res[name] = dict(ms_per_step=ms, us_per_launch=ms * 1e3 / L, GBps=gbs)
    peak = 6545.3
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for n, d in res.items():
        d["frac_of_measured_peak"] = d["GBps"] / peak
        d["frac_of_8TBps"] = d["GBps"] / 8000
    print(json.dumps(dict(geom=a.geom, bsz=B, T=T, layers=L, bytes_per_launch=bytes_per_launch, results=res)))


if __name__ == "__main__":
    main()
