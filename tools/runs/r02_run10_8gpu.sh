set +e
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/r02_8gpu_box.txt 2>&1
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02_bench_8gpu.json 2> gpurun_out/r02_bench_8gpu.err
echo "rc=$?" >> gpurun_out/r02_bench_8gpu.err
timeout 900 python bench.py --config engine --gpus 4 > gpurun_out/r02_bench_engine_4gpu.json 2> gpurun_out/r02_bench_engine_4gpu.err
tail -c 600 gpurun_out/r02_bench_8gpu.err
