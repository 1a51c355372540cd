#!/usr/bin/env python3
"""Largest difference between two pose CSV files (12 values per line): python tools/cmp_traj.py a.csv b.csv"""
import sys

import numpy as np


def load(p):
    rows = [[float(x) for x in l.replace(",", " ").split()][:12] for l in open(p) if l.strip()]
    return np.array(rows)


a, b = load(sys.argv[1]), load(sys.argv[2])
print({"poses": [len(a), len(b)], "max_abs_diff": float(np.abs(a - b).max()) if a.shape == b.shape else None,
       "max_abs_translation": float(np.abs(a[:, [3, 7, 11]]).max())})
