# one-GPU evidence run: ncu --set full of the hot kernels, the launch list of one timed epoch
set -x
python scripts/ncu_r02.py --big > gpurun_out/ncu_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'linear_p2_kernel|spmm_planned' -c 10 -f -o gpurun_out/r02_hot python scripts/ncu_r02.py --big > gpurun_out/ncu_hot.log 2>&1
echo "ncu hot rc $?"
python scripts/ncu_r02.py > gpurun_out/ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'knn_tc2' -c 2 -f -o gpurun_out/r02_knn python scripts/ncu_r02.py > gpurun_out/ncu_knn.log 2>&1
echo "ncu knn rc $?"
python bench.py --steps 1 --warmup 3 --epochs-per-step 1 --partition contiguous --no-e2e --no-agg-hbm --no-cpu-baseline > gpurun_out/r02_bench_launchlist_plain.json 2> gpurun_out/ncu_plain3.log && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_raw.csv python bench.py --steps 1 --warmup 3 --epochs-per-step 1 --partition contiguous --no-e2e --no-agg-hbm --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc $?"
ls -la gpurun_out/ | tail -12
