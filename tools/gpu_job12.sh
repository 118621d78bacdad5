#!/bin/bash
timeout 600 python -m pytest tests/test_primitives.py tests/test_solution.py -m gpu -q --tb=long 2>&1 | tail -60
