set +e
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_linear_gpu.py tests/test_model_gpu.py tests/test_wkv_chunk_gpu.py tests/test_engine_gpu.py -q -m gpu > gpurun_out/r02_suite_d.log 2>&1
echo "rc=$?" >> gpurun_out/r02_suite_d.log
timeout 900 python -m pytest tests/test_parity_full_gpu.py -q -s -k "config2 or config3" > gpurun_out/r02_parity_full_d.log 2>&1
echo "rc=$?" >> gpurun_out/r02_parity_full_d.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-engine --skip-cpu-baseline > gpurun_out/r02_bench_v3.json 2> gpurun_out/r02_bench_v3.err
B200_PAIR_BN=256 timeout 600 python bench.py --steps 30 --warmup 5 --no-engine --skip-cpu-baseline > gpurun_out/r02_bench_v3_bn256.json 2> gpurun_out/r02_bench_v3_bn256.err
timeout 600 python bench.py --config prefill --steps 3 --warmup 1 > gpurun_out/r02_bench_prefill_v4.json 2> gpurun_out/r02_bench_prefill_v4.err
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
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wkv7_chunk_kernel -s 40 -c 1 -f -o gpurun_out/r02_wkv_chunk_v4 python /tmp/prefill_once.py > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/r02_suite_d.log gpurun_out/r02_parity_full_d.log
