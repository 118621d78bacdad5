#!/bin/bash
# after the Lyapunov continue-from-state change (TbArgs grew: every kernel rebuilt): resume tests, full suite, default
# bench line, DRAM traffic of the rebuilt headline kernel
nvidia-smi -L
timeout 600 python -m pytest tests/test_lyap_resume.py tests/test_lyap_userjac.py -m gpu -x -q 2>&1 | tail -3
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_final.log 2>&1; tail -2 gpurun_out/r02_gputests_final.log
timeout 900 python bench.py > gpurun_out/r02_bench_c1_final.json 2> gpurun_out/r02_bench_c1_final.err; echo rc=$?; cut -c1-200 gpurun_out/r02_bench_c1_final.json
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:tb_ensemble -s 3 -c 1 --csv --log-file gpurun_out/r02_traffic_parity_final.csv \
    python bench.py --no-cpu-baseline --no-e2e --no-fast --steps 1 --warmup 3 > gpurun_out/r02_ncu_traffic_final.log 2>&1; echo rc=$?
tail -3 gpurun_out/r02_traffic_parity_final.csv
timeout 600 python bench.py --config 3 --no-cpu-baseline > gpurun_out/r02_bench_c3_check.json 2>/dev/null; cut -c1-160 gpurun_out/r02_bench_c3_check.json
