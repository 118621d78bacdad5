#!/bin/bash
# no-allocation TaylorN entry point: alone in a fresh process, then after the whole GPU suite
nvidia-smi -L
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -6
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -6
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
