set +e
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_parity_full_gpu.py > gpurun_out/r02_suite_f.log 2>&1
echo "rc=$?" >> gpurun_out/r02_suite_f.log
timeout 900 python -m pytest tests/test_parity_full_gpu.py -q -s -k "config2 or config3 or config1" > gpurun_out/r02_parity_full_f.log 2>&1
echo "rc=$?" >> gpurun_out/r02_parity_full_f.log
timeout 600 python bench.py --config prefill --steps 3 --warmup 1 > gpurun_out/r02_bench_prefill_v5.json 2> gpurun_out/r02_bench_prefill_v5.err
timeout 600 python bench.py --steps 20 --warmup 5 --no-engine --skip-cpu-baseline > gpurun_out/r02_bench_v4.json 2> gpurun_out/r02_bench_v4.err
timeout 200 python tools/bench_wkv.py --ref > gpurun_out/r02_wkv_vs_reference_kernel_v2.json 2>&1
timeout 200 python tools/bench_wkv.py --ref --geom 1.5B --bsz 64 >> gpurun_out/r02_wkv_vs_reference_kernel_v2.json 2>&1
cat > /tmp/prefill_once.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
import bench
geom, m = bench.build_model("7.2B", torch.device("cuda", 0))
toks = torch.randint(1, geom.vocab_size, (32, 1024), generator=torch.Generator().manual_seed(0)).to(torch.int32)
st = m.generate_zero_state(32)
m.prefill(toks, st, chunk_len=512, want_logits=False)
torch.cuda.synchronize()
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"(gemm|wkv7|ln_|embed|elapsed|tmix)" -c 1200 --csv --log-file gpurun_out/r02_prefill_launches_v2.csv python /tmp/prefill_once.py > gpurun_out/ncu1.log 2>&1
tail -n 3 gpurun_out/r02_suite_f.log gpurun_out/r02_parity_full_f.log
