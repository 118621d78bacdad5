#!/bin/bash
nvidia-smi -L
timeout 600 python -m pytest tests/test_tn.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --config 5a --ntraj 8192 --no-cpu-baseline --no-e2e --steps 2 --warmup 3 2>/dev/null | cut -c1-200
