N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "n$N rc $?"
python -c "import json; d=json.load(open('gpurun_out/r02_bench_n$N.json')); print('n$N', d['value'], d['config']['ms_per_epoch'], d['e2e']['value'], d['config']['setup_s'], d['graph_build']['setup_sections_s'], d['config']['gradient_averaging'], d['config']['cuda_graph_errors'])"
