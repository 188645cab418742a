set +e
mkdir -p gpurun_out
export CHIRRUP_B200_NO_BUILD=1
timeout 1200 python -m pytest tests/test_engine_gpu.py -m gpu -q -x > gpurun_out/r02_run22_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02_run22_tests.log
rm -f gpurun_out/r02_run22_ab.log
for i in 1 2; do
for b in 0 32; do
B200_ENGINE_BUCKET=$b timeout 900 python bench.py --config engine --skip-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('bucket $b', d['value'], e.get('steady_state_tok_per_s'), e.get('step_graph_hit_rate'), e.get('wall_s'))" >> gpurun_out/r02_run22_ab.log 2>&1
done
done
tail -n 4 gpurun_out/r02_run22_tests.log; cat gpurun_out/r02_run22_ab.log
