#!/bin/bash
# Scaling measurement on one 8-GPU box: bench.py at N = 1, 2, 4, 8 (ours), the reference arm at N = 1 and 8,
# correctness check on 8 real GPUs, and the larger BASELINE configs at N = 8.  Output: gpurun_out/scale/*.json
mkdir -p gpurun_out/scale
O=gpurun_out/scale
run() {  # name N extra-args...
    name=$1; N=$2; shift 2
    if [ "$N" = 1 ]; then
        timeout 200 python bench.py --gpus 1 "$@" > $O/$name.json 2> $O/$name.err
    else
        PORT=$((29600 + RANDOM % 300))
        timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > $O/$name.json 2> $O/$name.err
    fi
    echo "$name rc=$? $(tail -c 600 $O/$name.json | tr '\n' ' ' | cut -c1-330)"
}
run ref_er100k_n1 1 --impl reference
run ours_er100k_n1 1
run ours_er100k_n2 2
run ours_er100k_n4 4
run ours_er100k_n8 8
run ref_er100k_n8 8 --impl reference
PORT=$((29600 + RANDOM % 300))
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $PORT scripts/dist_check.py > $O/dist_check_n8.log 2>&1; echo "dist_check rc=$? $(tail -1 $O/dist_check_n8.log)"
run ours_rmat10m_n8 8 --config rmat10m
run ours_grid16m_n8 8 --config grid16m --steps 8 --warmup 3
run ours_kron26_n8 8 --config kron26 --steps 8 --warmup 3
