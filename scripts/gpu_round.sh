set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_pair_tier.py tests/test_gpu_search.py -x -q -m gpu > gpurun_out/pytest_pair.log 2>&1; echo "pair rc=$?"; tail -5 gpurun_out/pytest_pair.log
BIBFS_B200_LIB=build_ft/lib/libbibfs_b200.so timeout 200 python scripts/pair_finetrace.py er100k > gpurun_out/pair_finetrace.log 2>&1; echo "ft rc=$?"; head -24 gpurun_out/pair_finetrace.log
timeout 200 python scripts/latency_probe.py er100k 20 > gpurun_out/probe_er100k.log 2>&1; echo "probe rc=$?"; cat gpurun_out/probe_er100k.log | tail -8
