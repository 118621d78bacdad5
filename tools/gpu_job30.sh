#!/bin/bash
# round-2 final validation on one B200: full GPU suite, smoke, bench lines for configs 1-3 with the CPU leg and the
# parity sample, the reference arm, and two generator checks (events kernel with software pipelining, Lyapunov without)
nvidia-smi -L
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_final.log 2>&1; tail -3 gpurun_out/r02_gputests_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r02_bench_c1_final.json 2> gpurun_out/r02_bench_c1_final.err; echo rc=$?; cut -c1-200 gpurun_out/r02_bench_c1_final.json
for c in 2 3; do
timeout 900 python bench.py --config $c > gpurun_out/r02_bench_c${c}_final.json 2> gpurun_out/r02_bench_c${c}_final.err; echo rc=$?; cut -c1-200 gpurun_out/r02_bench_c${c}_final.json
done
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err; echo rc=$?; cut -c1-300 gpurun_out/r02_bench_reference_arm.json
echo "== config 2 with software pipelining"; timeout 600 python bench.py --config 2 --kflags 1207959552 --no-cpu-baseline --no-e2e 2>/dev/null | cut -c1-160
echo "== config 3 without software pipelining"; timeout 600 python bench.py --config 3 --kflags 32768 --no-cpu-baseline --no-e2e 2>/dev/null | cut -c1-160
