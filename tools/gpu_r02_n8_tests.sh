#!/bin/bash
# the multi-GPU parity tests on 8 GPUs, log kept
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -v 2>&1 | tail -14 | tee gpurun_out/r02_multi_gpu_pytest_n8.txt
