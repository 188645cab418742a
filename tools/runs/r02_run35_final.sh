set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1
echo "rc=$?" >> gpurun_out/r02_smoke_final.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err
timeout 900 python bench.py --config prefill --steps 3 --warmup 1 --skip-cpu-baseline > gpurun_out/r02_bench_prefill_final.json 2> gpurun_out/r02_bench_prefill_final.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"(gemm|wkv7|ln_|embed|elapsed|tmix|argmax)" -c 517 --csv --log-file gpurun_out/r02_prefill_launches_final.csv python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/ncu_p.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wkv7_chunk_kernel -s 4 -c 1 -f -o gpurun_out/r02_wkv_chunk_final python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/ncu_d.log 2>&1
tail -n 2 gpurun_out/r02_smoke_final.log
