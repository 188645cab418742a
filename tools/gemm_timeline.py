"""Debug aid: per-CTA timeline (globaltimer ns) of one GEMM launch: launch->prologue->first data->accumulators->partials parked
->cluster barrier->done."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200 import _lib, ops
L = _lib.lib()
L.b200_debug_gemm_probe.argtypes = [ctypes.c_void_p]; L.b200_debug_gemm_probe.restype = None
for (M, N, K, epi) in [(256, 4096, 4096, 0), (256, 4096, 16384, 4), (256, 16384, 4096, 3), (256, 4096, 128, 2), (256, 128, 4096, 1)]:
    x = torch.randn(M, K, device="cuda").half(); W = (torch.randn(N, K, device="cuda") / K ** 0.5).half()
    res = torch.randn(M, N, device="cuda").half() if epi == 4 else None
    bias = torch.randn(N, device="cuda").half() if epi in (1, 2) else None
    for _ in range(3): ops.linear(x, W, bias, epi, res)
    buf = torch.zeros(8 * 1024, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    L.b200_debug_gemm_probe(buf.data_ptr())
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); ops.linear(x, W, bias, epi, res); e.record()
    torch.cuda.synchronize()
    L.b200_debug_gemm_probe(0)
    t = buf.view(-1, 8).cpu()
    t = t[t[:, 0] > 0]
    t0 = t[:, 0].min()
    names = ["start", "prologue", "first_data", "accum_done", "parked", "cluster_sync", "epilogue_done"]
    print(f"M{M} N{N} K{K} epi{epi}: ctas {len(t)} event time {s.elapsed_time(e)*1e3:.1f} us")
    for i, n in enumerate(names):
        col = (t[:, i][t[:, i] > 0] - t0).float()
        if len(col): print(f"   {n:14s} min {col.min()/1e3:7.2f} med {col.median()/1e3:7.2f} max {col.max()/1e3:7.2f} us")
    w = t[:, 7][t[:, 7] >= (1 << 32)]
    mm = t[:, 7][(t[:, 7] > 0) & (t[:, 7] < (1 << 32))]
    if len(mm): print(f"   mma thread: of the issue cycles, the four tcgen05.mma alone {mm.float().median():.0f} cyc")
    if len(w):   # MMA-issuing thread of each leader CTA: cycles waiting for a full stage / issuing the 4 MMAs + commit, per K block
        kb = (K + 63) // 64
        wait, issue = (w >> 32).float(), (w & 0xffffffff).float()
        print(f"   mma thread: wait-for-data {wait.median():.0f} cyc, issue {issue.median():.0f} cyc in total over its K blocks "
              f"(K blocks per CTA <= {kb}); per-launch medians")
