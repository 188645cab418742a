set +e
mkdir -p gpurun_out
for n in 1 2 4; do
B200_MIG_STREAMS=$n timeout 600 python tools/bench_migration.py > gpurun_out/r02_migration_streams$n.json 2> gpurun_out/r02_migration_streams$n.err
done
B200_SLOT_COPY=kernel B200_SLOT_COPY_CTAS=148 timeout 600 python tools/bench_migration.py > gpurun_out/r02_migration_kernel148.json 2> gpurun_out/r02_migration_kernel148.err
B200_SLOT_COPY=kernel B200_SLOT_COPY_CTAS=148 B200_MIG_STREAMS=2 timeout 600 python tools/bench_migration.py > gpurun_out/r02_migration_kernel148_s2.json 2> gpurun_out/r02_migration_kernel148_s2.err
echo done
