set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 1500 python -m pytest tests/test_wkv_chunk_gpu.py tests/test_model_gpu.py tests/test_parity_full_gpu.py tests/test_engine_gpu.py -m gpu -q -x > gpurun_out/r02_run27_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02_run27_tests.log
rm -f gpurun_out/r02_run27_ab.log
for i in 1 2; do
for v in pers8 chunkB; do
CHIRRUP_B200_LIB=$PWD/tools/runs/_ab/lib_$v.so timeout 600 python bench.py --config prefill --steps 2 --warmup 1 --skip-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('prefill $v', d['ms_per_step'], d['value'], d['clocks']['sm_mhz'])" >> gpurun_out/r02_run27_ab.log 2>&1
done
done
tail -n 4 gpurun_out/r02_run27_tests.log; cat gpurun_out/r02_run27_ab.log
