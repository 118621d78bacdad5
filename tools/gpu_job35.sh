#!/bin/bash
# last check of the shipped tree: full GPU suite, smoke, default bench line (traffic stamp must match now)
nvidia-smi -L
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_final.log 2>&1; tail -2 gpurun_out/r02_gputests_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r02_bench_c1_final.json 2> gpurun_out/r02_bench_c1_final.err; echo rc=$?; cut -c1-160 gpurun_out/r02_bench_c1_final.json
timeout 600 python bench.py --config 4 > gpurun_out/r02_bench_c4_v4.json 2> gpurun_out/r02_bench_c4_v4.err; echo rc=$?; cut -c1-160 gpurun_out/r02_bench_c4_v4.json
TB_WIDE_NO_MUL2=1 timeout 600 python bench.py --config 4 --no-cpu-baseline 2>/dev/null | cut -c1-160
