for ov in 0 1; do
python bench.py --steps 10 --warmup 3 --partition contiguous --no-agg-hbm --no-cpu-baseline --no-e2e --overlap-eval $ov > gpurun_out/ab_ov_n1_$ov.json 2> gpurun_out/ab_ov_n1_$ov.err; python -c "import json; d=json.load(open('gpurun_out/ab_ov_n1_$ov.json')); print('n1 overlap=$ov', d['value'], d['config']['ms_per_epoch'], d['config']['final_train_loss'], d['config']['final_test_loss'], d['config']['cuda_graphs'])"
done
