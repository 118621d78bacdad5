#!/bin/bash
timeout 900 python -m pytest tests/test_wide.py tests/test_primitives.py -m gpu -x -q 2>&1 | tail -4
python tools/bench_configs.py manyspin 2>&1 | cut -c1-600
python tools/bench_configs.py manyspin --fast 2>&1 | cut -c1-300
