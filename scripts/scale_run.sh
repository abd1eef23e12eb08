cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-2}
export BIBFS_BENCH_VERBOSE=1
if [ "$2" = "check" ]; then timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 scripts/dist_check.py > gpurun_out/dist_check_n$N.log 2>&1; echo "dist_check rc=$?"; tail -3 gpurun_out/dist_check_n$N.log; fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 20 --warmup 5 --others "${3:-}" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; python -c "
import json,sys; d=json.loads([l for l in open('gpurun_out/bench_n$N.json') if l.startswith('{')][-1]); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','vs_baseline','n_gpus')}, 'e2e', d['e2e']['ms_per_step'], d['search_ms'], d.get('other_configs'))"; grep "rank" gpurun_out/bench_n$N.err | head -20
