"""BASELINE.md §4, last bullet: the "reference on B200" step time — the oracle (oracle/rwkv7_oracle.py: the reference's ATen op
sequence, torch/cuBLAS GEMMs, ~60 launches per layer) with the reference's OWN CUDA WKV kernel (oracle/_ref/libref_wkv7.so) on
the same box, timed at BASELINE configs[1] (1.5B, bsz 64) and configs[2] (7.2B, bsz 256), next to our step.  Wall clock around
synchronised loops (the reference kernel launches on the legacy stream and the wrapper synchronises, exactly like its torch op
would serialise with the default stream).  TEST/BENCH INFRASTRUCTURE: this is the only tool that times oracle/ on the GPU.
Usage (GPU box): python tools/bench_oracle_gpu.py [--configs 1.5B:64,7.2B:256] [--steps 10]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200.rwkv7 import RWKV_x070, make_args
from chirrup_b200.synth import GEOMETRIES, make_checkpoint
from oracle import rwkv7_oracle as O


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1.5B:64,7.2B:256")
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    for cfg in a.configs.split(","):
        name, bsz = cfg.split(":")
        B = int(bsz)
        geom = GEOMETRIES[name]
        raw = make_checkpoint(geom, seed=0, recipe="trained_like", device="cuda", dtype=torch.float16)
        m = RWKV_x070(make_args(), state_dict={k: v.clone() for k, v in raw.items()})
        z = O.transform_checkpoint(raw, device="cuda")
        og = O.Geometry(z)
        del raw
        toks = torch.randint(1, geom.vocab_size, (B, 1), generator=torch.Generator().manual_seed(0)).tolist()
        sa, sb = m.generate_zero_state(B), O.zero_state(og, B, "cuda")
        res = {}
        for label, fn in (("oracle_torch_cublas_plus_reference_wkv_kernel", lambda: O.forward_seq_batch(z, og, toks, sb, wkv="ref_cuda")),
                          ("ours_public_api", lambda: m.forward_seq_batch_seperate(toks, sa))):
            for _ in range(16):
                fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(a.steps):
                fn()
            torch.cuda.synchronize()
            res[label] = {"ms_per_step": 1e3 * (time.perf_counter() - t0) / a.steps}
            res[label]["tok_per_s"] = B / res[label]["ms_per_step"] * 1e3
        res["speedup"] = res["oracle_torch_cublas_plus_reference_wkv_kernel"]["ms_per_step"] / res["ours_public_api"]["ms_per_step"]
        print(json.dumps({"config": f"{name} (L{geom.n_layer}, C{geom.n_embd}) decode, bsz {B}, T=1", **res}), flush=True)
        del m, z, sa, sb
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
