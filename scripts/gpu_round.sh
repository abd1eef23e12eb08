cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 10 --warmup 5 --others "rmat22_batch64,kron26_batch64" > gpurun_out/bench_batch_n1.json 2> gpurun_out/bench_batch_n1.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_batch_n1.json')); print(d['ms_per_step']); print(json.dumps(d['other_configs'], indent=1))"; tail -5 gpurun_out/bench_batch_n1.err
