#!/bin/bash
# quick GPU iteration: parity tests (optionally filtered) + kernel timing matrix
tag=${1:-q}; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS} > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -15 $out/pytest_gpu.log
timeout 900 python tools/kbench.py "$@" > $out/kbench.jsonl 2> $out/kbench.err; echo "kbench exit $?"
cat $out/kbench.jsonl; tail -5 $out/kbench.err
