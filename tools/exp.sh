#!/bin/bash
set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/kernel_times.py --frames 1000 --batch 250 --tag select_hybrid
