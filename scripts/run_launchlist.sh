# launch list of ONE timed epoch of the final code (the bench command first runs without ncu and must exit 0)
python bench.py --steps 1 --warmup 3 --epochs-per-step 1 --no-e2e --no-agg-hbm --no-cpu-baseline > gpurun_out/r02_bench_launchlist_plain.json 2> gpurun_out/ncu_plain3.log && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_raw.csv python bench.py --steps 1 --warmup 3 --epochs-per-step 1 --no-e2e --no-agg-hbm --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc $?"
wc -l gpurun_out/r02_launches_raw.csv
