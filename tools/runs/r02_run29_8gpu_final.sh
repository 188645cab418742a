set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
for n in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 --skip-cpu-baseline > gpurun_out/r02_bench_${n}gpu_final.json 2> gpurun_out/r02_bench_${n}gpu_final.err
tail -n 1 gpurun_out/r02_bench_${n}gpu_final.json | cut -c1-200
done
