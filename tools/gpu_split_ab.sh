#!/bin/bash
# parity tests of the IEKF path on the new build, then A/B of LIO_TILE_SPLIT on the bench workload
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_lio.py -m gpu -x -q 2>&1 | tail -4
bash tools/gpu_ab.sh "LIO_TILE_SPLIT=0" "LIO_TILE_SPLIT=8" "LIO_TILE_SPLIT=4" "LIO_TILE_SPLIT=0" "LIO_TILE_SPLIT=8"
