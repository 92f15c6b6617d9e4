port=29600
for ov in 0 1; do
port=$((port+1))
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 20 --warmup 3 --overlap-eval $ov > gpurun_out/r02_bench_n8_ov$ov.json 2> gpurun_out/r02_bench_n8_ov$ov.err; echo "n8 ov=$ov rc $?"
python -c "import json; d=json.load(open('gpurun_out/r02_bench_n8_ov$ov.json')); print('n8 overlap=$ov', d['value'], d['config']['ms_per_epoch'], d['e2e']['value'], d['config']['setup_s'], d['graph_build']['setup_sections_s'], d['config']['gradient_averaging'], d['config']['cuda_graph_errors'])"
done
