python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "n1 rc $?"
python -c "import json; d=json.load(open('gpurun_out/r02_bench_n1.json')); print(d['value'], d['config']['ms_per_epoch'], d['e2e']['value'], d['e2e']['h2d_gbs_achieved'], d['cpu_baseline']['value'], d['roofline']['frac'], d['roofline']['traffic'], d['config']['setup_s'])"
python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err; echo "ref rc $?"
python -c "import json; d=json.load(open('gpurun_out/r02_bench_reference.json')); print(d['value'], d['ms_per_step'], d['config']['epoch_fraction_per_step']); print(d['cpu_baseline']['sample'])"
