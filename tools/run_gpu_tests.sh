#!/bin/bash
# usage (on the GPU box, from the repo root): bash tools/run_gpu_tests.sh
timeout 1700 python -m pytest tests -m gpu -q -rA 2>&1 | grep -v "^ERROR in\|^$" | tail -80
