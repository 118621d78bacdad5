#!/bin/bash
# bench.py (new): config 1 full line, configs 2 and 3, then the reference arm
python bench.py > gpurun_out/r02_bench_c1.json 2> gpurun_out/r02_bench_c1.err; echo rc=$?; cut -c1-1800 gpurun_out/r02_bench_c1.json; tail -3 gpurun_out/r02_bench_c1.err
python bench.py --config 2 > gpurun_out/r02_bench_c2.json 2> gpurun_out/r02_bench_c2.err; echo rc=$?; cut -c1-1800 gpurun_out/r02_bench_c2.json; tail -3 gpurun_out/r02_bench_c2.err
python bench.py --config 3 > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err; echo rc=$?; cut -c1-1800 gpurun_out/r02_bench_c3.json; tail -3 gpurun_out/r02_bench_c3.err
