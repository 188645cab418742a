set +e
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -q -m gpu > gpurun_out/r02_engine_gpu_2gpu_b.log 2>&1
echo "rc=$?" >> gpurun_out/r02_engine_gpu_2gpu_b.log
for n in 1 2 4; do
B200_MIG_STREAMS=$n timeout 600 python tools/bench_migration.py > gpurun_out/r02_migration_batched_streams$n.json 2> gpurun_out/r02_migration_batched_streams$n.err
done
B200_SLOT_COPY=kernel B200_SLOT_COPY_CTAS=148 timeout 600 python tools/bench_migration.py > gpurun_out/r02_migration_batched_kernel148.json 2> gpurun_out/r02_migration_batched_kernel148.err
tail -n 3 gpurun_out/r02_engine_gpu_2gpu_b.log
