#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_rowsplit.py -m gpu -x -q > $O/pytest_rowsplit.log 2>&1; echo "pytest rowsplit rc=$?"; tail -3 $O/pytest_rowsplit.log
