#!/bin/bash
timeout 900 python -m pytest tests/test_bf_gpu.py -x -q 2>&1 | tail -15
python profiles/bf_sizes_timing.py 2>&1 | tail -1
