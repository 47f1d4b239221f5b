#!/bin/bash
# Device-loop validation pass: lio parity tests, bench without the loop section, launch list, ncu full capture of the
# second-tier kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lio.py tests/test_abi.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-loop > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-2500 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
LIO_IEKF_MODE=host timeout 600 python bench.py --no-loop --no-cpu-baseline > gpurun_out/bench_hostloop.json 2> gpurun_out/bench_hostloop.err; echo "bench(host loop) exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_hostloop.json'));print('host loop', d['value'], d['e2e']['value'], d['kernel_ms'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_plane_reduce|k_knn_left|k_ds_commit|k_ds_scan|k_cells_ensure|k_insert_plain' -s 30 -c 14 -o gpurun_out/tier2_full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-loop > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
