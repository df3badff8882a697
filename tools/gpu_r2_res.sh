#!/bin/bash
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python tools/conv_probe.py 4 | tail -3
PH=128 timeout 300 python tools/conv_probe.py 4 | tail -1
