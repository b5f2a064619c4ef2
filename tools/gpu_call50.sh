#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 100 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke50.log 2>&1
timeout 150 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/t50.log 2>&1
echo "rc=$?" >> gpurun_out/t50.log
