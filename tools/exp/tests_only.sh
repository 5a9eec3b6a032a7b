#!/bin/bash
# gpurun payload: GPU tests selected by the arguments (pytest syntax), full tracebacks of the first failure
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest "$@" -m gpu -x -q > gpurun_out/pytest_sel.log 2>&1; echo "rc=$?"
tail -40 gpurun_out/pytest_sel.log
