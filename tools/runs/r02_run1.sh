set +e
mkdir -p gpurun_out
(nproc; free -g; nvidia-smi --query-gpu=name,memory.total --format=csv) > gpurun_out/r02_box.txt 2>&1
timeout 1200 python -m pytest tests/test_parity_full_gpu.py -q -s > gpurun_out/r02_parity_full.log 2>&1
echo "parity rc=$?" >> gpurun_out/r02_parity_full.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_parity_full_gpu.py > gpurun_out/r02_gpu_suite.log 2>&1
echo "suite rc=$?" >> gpurun_out/r02_gpu_suite.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_v1.json 2> gpurun_out/r02_bench_v1.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_v1.json 2> gpurun_out/r02_bench_ref_v1.err
timeout 400 python tools/bench_gemm.py --cublas --json gpurun_out/r02_gemm_vs_cublas.json > gpurun_out/r02_gemm_vs_cublas.log 2>&1
timeout 200 python tools/bench_wkv.py --ref > gpurun_out/r02_wkv_vs_reference_kernel.json 2>&1
timeout 200 python tools/bench_wkv.py --ref --geom 1.5B --bsz 64 >> gpurun_out/r02_wkv_vs_reference_kernel.json 2>&1
timeout 400 python tools/bench_oracle_gpu.py > gpurun_out/r02_oracle_gpu_step.json 2>&1
tail -3 gpurun_out/r02_parity_full.log gpurun_out/r02_gpu_suite.log
