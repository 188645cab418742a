set +e
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/r02_2gpu_box.txt 2>&1
timeout 900 python -m pytest tests/test_engine_gpu.py -q -m gpu > gpurun_out/r02_engine_gpu_2gpu.log 2>&1
echo "rc=$?" >> gpurun_out/r02_engine_gpu_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err
timeout 900 python bench.py --config engine --gpus 2 > gpurun_out/r02_bench_engine_2gpu.json 2> gpurun_out/r02_bench_engine_2gpu.err
timeout 900 python bench.py --config engine --gpus 1 > gpurun_out/r02_bench_engine_1gpu.json 2> gpurun_out/r02_bench_engine_1gpu.err
timeout 900 python tools/bench_migration.py > gpurun_out/r02_migration_copy_engine.json 2> gpurun_out/r02_migration_copy_engine.err
B200_SLOT_COPY=kernel timeout 900 python tools/bench_migration.py > gpurun_out/r02_migration_kernel64.json 2> gpurun_out/r02_migration_kernel64.err
B200_SLOT_COPY=kernel B200_SLOT_COPY_CTAS=32 timeout 900 python tools/bench_migration.py > gpurun_out/r02_migration_kernel32.json 2> gpurun_out/r02_migration_kernel32.err
tail -n 3 gpurun_out/r02_engine_gpu_2gpu.log
