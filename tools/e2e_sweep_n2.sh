#!/bin/bash
# e2e sweep at N=2
run(){ tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 2 --steps 5 --warmup 3 --no-configs --no-cpu > gpurun_out/r02_sw_$tag.json 2> gpurun_out/r02_sw_$tag.err; }
run base A=1
run nw2 P2PT_BENCH_NW=2
run nw4 P2PT_BENCH_NW=4
run nb16 P2PT_BENCH_NB=16
run c32 P2PT_BENCH_CHUNK_MB=32
run nw2c32 P2PT_BENCH_NW=2 P2PT_BENCH_CHUNK_MB=32
