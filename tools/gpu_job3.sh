#!/bin/bash
# TMEM microbench + first TMEM-tier kernels: parity check vs the oracle and timing
timeout 120 ./tools/tmem_bench > gpurun_out/r02_tmem_bench.txt 2>&1; tail -45 gpurun_out/r02_tmem_bench.txt
timeout 900 python tools/tmem_sweep.py > gpurun_out/r02_tmem_sweep.txt 2>&1; cat gpurun_out/r02_tmem_sweep.txt
