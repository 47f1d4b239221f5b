#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --no-loop --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print(d['value'], d['e2e']['value'], d['ms_per_step'], d['kernel_ms'], d['roofline']['frac'])"; tail -3 gpurun_out/bench.err
timeout 900 python bench.py --workload sequence --seq-scans 60 > gpurun_out/bench_seq.json 2> gpurun_out/bench_seq.err; echo "sequence exit $?"; cut -c1-1500 gpurun_out/bench_seq.json; tail -3 gpurun_out/bench_seq.err
