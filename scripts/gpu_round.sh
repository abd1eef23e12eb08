set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
BIBFS_B200_LIB=build_ft/lib/libbibfs_b200.so timeout 200 python scripts/pair_finetrace.py er100k > gpurun_out/pair_finetrace.log 2>&1; echo "ft rc=$?"; head -50 gpurun_out/pair_finetrace.log
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
