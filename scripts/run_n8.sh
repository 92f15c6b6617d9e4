# 8-GPU box: A/B of the all-reduce cut, then the full default bench at N = 8 and N = 4
port=29540
for cut in none last_wave; do
port=$((port+1))
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 20 --warmup 3 --partition contiguous --no-e2e --no-agg-hbm --dp-cut $cut > gpurun_out/ab_n8_$cut.json 2> gpurun_out/ab_n8_$cut.err; echo "$cut rc $?"; python -c "import json; d=json.load(open('gpurun_out/ab_n8_$cut.json')); print('n8 $cut', d['value'], d['config']['ms_per_epoch'])"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29550 bench.py --gpus 8 --steps 20 --warmup 3 --dp-cut none > gpurun_out/bench_full_n8.json 2> gpurun_out/bench_full_n8.err; echo "full n8 rc $?"; tail -3 gpurun_out/bench_full_n8.err; python -c "import json; d=json.load(open('gpurun_out/bench_full_n8.json')); print('n8 full', d['value'], d['config']['ms_per_epoch'], d['config']['setup_s'], d['graph_build']['setup_sections_s'], d['e2e']['value'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 4 --steps 20 --warmup 3 --dp-cut none > gpurun_out/bench_full_n4.json 2> gpurun_out/bench_full_n4.err; echo "full n4 rc $?"; python -c "import json; d=json.load(open('gpurun_out/bench_full_n4.json')); print('n4 full', d['value'], d['config']['ms_per_epoch'], d['config']['setup_s'], d['e2e']['value'])"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29552 scripts/probe_symm.py 2>&1 | grep "nccl_us"
