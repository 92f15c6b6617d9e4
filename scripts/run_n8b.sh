timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29581 scripts/probe_nvls.py 2>&1 | grep world
port=29582
for ar in p2p nvls nccl; do
port=$((port+1))
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 20 --warmup 3 --partition contiguous --no-e2e --no-agg-hbm --dp-cut none --allreduce $ar > gpurun_out/ab2_n8_$ar.json 2> gpurun_out/ab2_n8_$ar.err; echo "$ar rc $?"; python -c "import json; d=json.load(open('gpurun_out/ab2_n8_$ar.json')); print('n8 $ar', d['value'], d['config']['ms_per_epoch'], d['config']['cuda_graphs'], d['config']['cuda_graph_errors'], d['config']['gradient_averaging'])"
done
