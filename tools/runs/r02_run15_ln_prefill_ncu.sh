set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 600 python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/r02_run15_prefill.json 2> gpurun_out/r02_run15_prefill.err
echo "rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ln_shift_mix_kernel -s 40 -c 4 -f -o gpurun_out/r02_ln_prefill python bench.py --config prefill --steps 1 --warmup 1 --skip-cpu-baseline > gpurun_out/ncu_ln.log 2>&1
echo "rc=$?"
tail -n 2 gpurun_out/r02_run15_prefill.json | cut -c1-600
