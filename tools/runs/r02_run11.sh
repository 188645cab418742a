set +e
mkdir -p gpurun_out
for mb in 0 24 48 96; do
B200_L2_PREFETCH_MB=$mb timeout 600 python bench.py --steps 40 --warmup 10 --no-engine --skip-cpu-baseline > gpurun_out/r02_bench_pf$mb.json 2> gpurun_out/r02_bench_pf$mb.err
done
B200_L2_PREFETCH_MB=0 timeout 600 python bench.py --steps 40 --warmup 10 --no-engine --skip-cpu-baseline > gpurun_out/r02_bench_pf0_b.json 2> gpurun_out/r02_bench_pf0_b.err
timeout 900 python -m pytest tests/test_model_gpu.py -q -m gpu > gpurun_out/r02_suite_e.log 2>&1
echo "rc=$?" >> gpurun_out/r02_suite_e.log
tail -n 3 gpurun_out/r02_suite_e.log
