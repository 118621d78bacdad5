#!/bin/bash
timeout 900 python -m pytest tests/test_tn.py -m gpu -x -q 2>&1 | tail -3
python tools/tn_probe.py 148 0.5
python tools/tn_probe.py 2048 6.283185307179586
python tools/tn_probe.py 16384 20 2 8
