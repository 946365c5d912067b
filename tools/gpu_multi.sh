#!/bin/bash
# Multi-GPU session: bench.py and the sharded prover under torchrun at N ranks.
# usage: gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_multi.sh TAG N'
TAG=${1:-r01m}; N=${2:-2}
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/${TAG}_smi.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "bench N=$N exit $?"
tail -c 600 $O/${TAG}_bench_n$N.json
timeout 900 $TR tools/bench_sharded_prove.py --log-n ${SIZES:-16 20 22} --check > $O/${TAG}_prove_n$N.json 2> $O/${TAG}_prove_n$N.err; echo "sharded prove N=$N exit $?"
cat $O/${TAG}_prove_n$N.json
tail -5 $O/${TAG}_prove_n$N.err
