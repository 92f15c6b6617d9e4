python -m pytest tests/test_gpu_trainers.py -m gpu -q -x -k "adam" 2>&1 | tail -2
for o in own torch; do
python bench.py --steps 10 --warmup 3 --partition contiguous --no-agg-hbm --no-cpu-baseline --no-e2e --optimizer $o > gpurun_out/ab_opt_$o.json 2> gpurun_out/ab_opt_$o.err; python -c "import json; d=json.load(open('gpurun_out/ab_opt_$o.json')); print('$o', d['value'], d['config']['ms_per_epoch'], d['config']['final_train_loss'], d['config']['final_test_loss'])"
done
