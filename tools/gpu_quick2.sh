#!/bin/bash
out=gpurun_out/${1:-quick2}; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pseudo_adaptive.py tests/test_gpu_pa_remap.py tests/test_zz_no_alloc.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -2 $out/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
