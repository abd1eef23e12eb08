cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 300 python scripts/batch_probe.py rmat20 2>&1 | tail -5
timeout 300 python scripts/batch_probe.py er100k 2>&1 | tail -5
