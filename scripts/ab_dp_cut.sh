# A/B timings: where the gradient all-reduce is cut (2 GPUs) and side-stream forks on / off (1 GPU); fast set-up
port=29520
for cut in none last_wave every_wave; do
port=$((port+1))
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --steps 10 --warmup 3 --partition contiguous --no-e2e --no-agg-hbm --dp-cut $cut > gpurun_out/ab_n2_$cut.json 2> gpurun_out/ab_n2_$cut.err; echo "$cut rc $?"; python -c "import json; d=json.load(open('gpurun_out/ab_n2_$cut.json')); print('$cut', d['value'], d['config']['ms_per_epoch'])"
done
for f in 1 0; do
GNNB_FORK=$f python bench.py --steps 5 --warmup 3 --partition contiguous --no-e2e --no-agg-hbm --no-cpu-baseline > gpurun_out/ab_n1_fork$f.json 2> gpurun_out/ab_n1_fork$f.err; python -c "import json; d=json.load(open('gpurun_out/ab_n1_fork$f.json')); print('n1 fork=$f', d['value'], d['config']['ms_per_epoch'])"
done
