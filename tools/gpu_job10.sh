#!/bin/bash
timeout 900 python -m pytest tests/test_wide.py -m gpu -x -q 2>&1 | tail -3
python tools/bench_configs.py manyspin 2>&1 | cut -c1-420
TB_WIDE_LC_GLOBAL=1 python tools/bench_configs.py manyspin 2>&1 | cut -c1-300
