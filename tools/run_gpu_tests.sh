#!/bin/bash
# usage (on the GPU box, from the repo root): bash tools/run_gpu_tests.sh
timeout 1500 python -m pytest tests -m gpu -x -q -rA 2>&1 | tail -60
