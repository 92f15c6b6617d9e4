python -m pytest tests -m gpu -q > gpurun_out/r2_tests_final.log 2>&1; echo "tests exit $?"; grep -E "passed|failed|^FAILED" gpurun_out/r2_tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "n1 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_n1.json')); print(d['value'], d['config']['ms_per_epoch'], d['e2e']['value'], d['cpu_baseline']['value'], d['config']['setup_s'], d['roofline']['frac'], d['gpu_launches'])"
