#!/bin/bash
out=gpurun_out/${1:-r02bc}; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_pageable.py -m gpu -x -q > $out/pytest_pageable.log 2>&1
echo "pytest pageable exit $?"; tail -4 $out/pytest_pageable.log
timeout 1200 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -2 $out/pytest_gpu.log
