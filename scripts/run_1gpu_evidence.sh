python -m pytest tests -m gpu -q --durations=5 > gpurun_out/r2_tests5.log 2>&1; echo "tests exit $?"; grep -E "passed|failed|^FAILED" gpurun_out/r2_tests5.log
bash scripts/run_ncu.sh > gpurun_out/run_ncu.log 2>&1; grep -E "rc " gpurun_out/run_ncu.log
for c in C1 C2 C3 C4; do python bench.py --config $c --steps 2 > gpurun_out/r02_bench_$c.json 2> gpurun_out/r02_bench_$c.err; echo "$c rc $?"; python -c "import json; d=json.load(open('gpurun_out/r02_bench_$c.json')); print('$c', d['value'], d['cpu_baseline']['value'], d['e2e']['value'], d['config']['graph_build_s'])"; done
