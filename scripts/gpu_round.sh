cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 200 python scripts/latency_probe.py er100k 20 > gpurun_out/probe_er100k.log 2>&1; echo "probe rc=$?"; cat gpurun_out/probe_er100k.log | tail -6
BIBFS_B200_LIB=build_ft/lib/libbibfs_b200.so timeout 200 python scripts/pair_finetrace.py er100k > gpurun_out/pair_finetrace.log 2>&1; echo "ft rc=$?"; grep -E "^cold|^hot" gpurun_out/pair_finetrace.log
