python -m pytest tests -m gpu -q > gpurun_out/r2_tests_final.log 2>&1; echo "tests exit $?"; grep -E "passed|failed|^FAILED" gpurun_out/r2_tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "n1 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_n1.json')); print(d['value'], d['config']['ms_per_epoch'], d['e2e']['value'], d['cpu_baseline']['value'], d['config']['setup_s'], d['graph_build']['setup_sections_s'])
for k in d['graph_build']['knn']: print(k['call'], k['calls'], round(k['avg_ms'],2), round(k['tops']), round(k['frac_of_4x_bf16_peak'],3))"
python scripts/ncu_fp4.py > gpurun_out/ncu_fp4_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:knn_tc2 -c 2 -f -o gpurun_out/r02_knn_fp4 python scripts/ncu_fp4.py > gpurun_out/ncu_fp4.log 2>&1; echo ncu rc $?
