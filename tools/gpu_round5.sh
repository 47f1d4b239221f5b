#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_imu.py -m gpu -x -q > gpurun_out/pytest_imu.log 2>&1; echo "pytest imu exit $?"; tail -8 gpurun_out/pytest_imu.log
timeout 900 python tools/tile_sweep.py > gpurun_out/tile_sweep.log 2>&1; echo "sweep exit $?"; cat gpurun_out/tile_sweep.log | tail -20
timeout 600 ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -c 400 --csv --log-file gpurun_out/launches_warm.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_launch.log 2>&1; echo "ncu warm launches exit $?"
