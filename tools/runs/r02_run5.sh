set +e
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_gpu_suite_full.log 2>&1
echo "rc=$?" >> gpurun_out/r02_gpu_suite_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1
echo "rc=$?" >> gpurun_out/r02_smoke.log
timeout 600 python bench.py --config prefill --steps 3 --warmup 1 > gpurun_out/r02_bench_prefill_v3.json 2> gpurun_out/r02_bench_prefill_v3.err
timeout 600 python bench.py --config engine --gpus 1 > gpurun_out/r02_bench_engine_1gpu_v2.json 2> gpurun_out/r02_bench_engine_1gpu_v2.err
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
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wkv7_chunk_kernel -s 40 -c 1 -f -o gpurun_out/r02_wkv_chunk_v3 python /tmp/prefill_once.py > gpurun_out/ncu2.log 2>&1
tail -n 4 gpurun_out/r02_gpu_suite_full.log gpurun_out/r02_smoke.log
