#!/usr/bin/env python
"""Bit-for-bit comparison of two .npz dumps (tools/dense_probe.py --dump).  usage: python tools/npz_equal.py a.npz b.npz"""
import sys

import numpy as np

a, b = np.load(sys.argv[1]), np.load(sys.argv[2])
ok = sorted(a.files) == sorted(b.files) and all(
    a[k].shape == b[k].shape and np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)) for k in a.files)
print("equal" if ok else "DIFFERENT", sys.argv[1], sys.argv[2])
sys.exit(0 if ok else 1)
