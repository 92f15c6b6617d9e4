timeout 300 python -m pytest tests/test_gpu_dp_nccl.py -m gpu -q -x > gpurun_out/r2_nccl_test2.log 2>&1; echo "nccl test exit $?"; tail -15 gpurun_out/r2_nccl_test2.log
port=29560
for ar in p2p nvls nccl; do
port=$((port+1))
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --steps 10 --warmup 3 --partition contiguous --no-e2e --no-agg-hbm --dp-cut none --allreduce $ar > gpurun_out/ab2_n2_$ar.json 2> gpurun_out/ab2_n2_$ar.err; echo "$ar rc $?"; python -c "import json; d=json.load(open('gpurun_out/ab2_n2_$ar.json')); print('n2 $ar', d['value'], d['config']['ms_per_epoch'], d['config']['cuda_graphs'], d['config']['cuda_graph_errors'], d['config']['gradient_averaging'])"
done
