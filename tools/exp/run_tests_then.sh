#!/bin/bash
# gpurun payload: the named GPU tests, smoke(), then another payload script with its arguments
# usage: run_tests_then.sh "<pytest paths>" <script> [args...]
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest $1 -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/pytest_sel.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
shift
bash "$@"
