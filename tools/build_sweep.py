#!/usr/bin/env python
"""Builds the map-kernel tuning variants swept by tools/kbench.py."""
import itertools, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "psana-ray_b200"))
import build
V = []
for thr, unr, minb, ng in [(256, 4, 3, 2), (256, 8, 2, 2), (128, 4, 6, 2), (512, 4, 1, 2),
                           (256, 4, 4, 1), (256, 8, 4, 1), (256, 4, 6, 1), (128, 4, 8, 1), (128, 8, 8, 1), (512, 4, 2, 1), (512, 8, 2, 1)]:
    V.append((f"t{thr}_u{unr}_b{minb}_g{ng}", dict(PSR_MAP_THREADS=thr, PSR_MAP_UNROLL=unr, PSR_MAP_MINB=minb, PSR_MAP_GROUPS=ng)))
for name, d in V:
    print(build.build_variant(name, d))
