#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > $O/r02l_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02l_pytest.log
tail -3 $O/r02l_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02l_smoke.log 2>&1; tail -1 $O/r02l_smoke.log
