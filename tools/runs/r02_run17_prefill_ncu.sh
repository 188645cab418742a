set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 600 python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/b17.log 2>&1
echo "rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"(gemm|wkv7|ln_|embed|elapsed|tmix|argmax)" -c 517 --csv --log-file gpurun_out/r02_prefill_launches_v6.csv python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/ncu17a.log 2>&1
echo "rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_gemm_persist_kernel -s 20 -c 5 -f -o gpurun_out/r02_persist_gemm python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/ncu17b.log 2>&1
echo "rc=$?"
