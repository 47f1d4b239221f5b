#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_loop.py -m gpu -x -q > gpurun_out/pytest_gpu_loop.log 2>&1; echo "pytest loop exit $?"; tail -3 gpurun_out/pytest_gpu_loop.log
timeout 900 python tools/tile_sweep.py > gpurun_out/tile_sweep2.log 2>&1; echo "sweep exit $?"; cat gpurun_out/tile_sweep2.log | cut -c1-200
