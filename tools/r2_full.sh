#!/bin/bash
# whole GPU suite + smoke (quick confirmation between changes)
O=gpurun_out; T=${1:-r02}
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -6 $O/pytest_gpu_$T.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$T.log 2>&1; echo "smoke exit $?"; tail -1 $O/smoke_$T.log
