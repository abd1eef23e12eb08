set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" 
timeout 300 python -m pytest tests/test_gpu_pair_tier.py -x -q -m gpu > gpurun_out/pytest_pair.log 2>&1; echo "pair rc=$?"; tail -15 gpurun_out/pytest_pair.log
timeout 200 python scripts/latency_probe.py er100k 20 > gpurun_out/probe_er100k.log 2>&1; echo "probe rc=$?"; cat gpurun_out/probe_er100k.log | tail -8
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_n1.json; tail -5 gpurun_out/bench_n1.err
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
