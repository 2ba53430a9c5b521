#!/bin/bash
# bench.py at N = $GPUS ranks (torchrun, one rank per GPU), own arm only
N=${GPUS:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N ${BENCH_ARGS:-} > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$?"; tail -n 5 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
