set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_gpu_suite_2gpu_final.log 2>&1
echo "rc=$?" >> gpurun_out/r02_gpu_suite_2gpu_final.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_2gpu_final.json 2> gpurun_out/r02_bench_2gpu_final.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_2gpu_final.json 2> gpurun_out/r02_bench_reference_2gpu_final.err
timeout 600 python bench.py --config migration > gpurun_out/r02_migration_final.json 2> gpurun_out/r02_migration_final.err
tail -n 3 gpurun_out/r02_gpu_suite_2gpu_final.log; tail -n 1 gpurun_out/r02_bench_2gpu_final.json | cut -c1-300; tail -n 1 gpurun_out/r02_bench_reference_2gpu_final.json | cut -c1-300; tail -n 1 gpurun_out/r02_migration_final.json | cut -c1-400
