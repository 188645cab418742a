set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_noalloc_alone.log 2>&1
echo "rc=$?" >> gpurun_out/r02_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r02_gpu_suite_with_noalloc.log 2>&1
echo "rc=$?" >> gpurun_out/r02_gpu_suite_with_noalloc.log
tail -n 15 gpurun_out/r02_noalloc_alone.log; tail -n 4 gpurun_out/r02_gpu_suite_with_noalloc.log
