#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/r02k_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02k_pytest.log
tail -15 $O/r02k_pytest.log | cut -c1-400
