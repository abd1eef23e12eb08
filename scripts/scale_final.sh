#!/bin/bash
# final check of the round on an 8-GPU box: ours at N = 2, 4, 8 (er100k) and the correctness check on 8 GPUs
mkdir -p gpurun_out/scale_final; O=gpurun_out/scale_final
for N in 2 4 8; do
  PORT=$((29600 + RANDOM % 300))
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N > $O/ours_er100k_n$N.json 2> $O/ours_er100k_n$N.err
  echo "N=$N rc=$? $(tail -1 $O/ours_er100k_n$N.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4), round(d["e2e"]["ms_per_step"],4), d["replicated_graph_throughput"]["queries_per_s"] if d.get("replicated_graph_throughput") else None)')"
done
PORT=$((29600 + RANDOM % 300))
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $PORT scripts/dist_check.py > $O/dist_check_n8.log 2>&1; echo "dist_check rc=$? $(tail -1 $O/dist_check_n8.log)"
