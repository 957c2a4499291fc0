#!/usr/bin/env python
"""Builds variants of libgsrast.so (extra -D defines) under gaussian_splatting_b200/variants/ for A/B runs on the
GPU box (profiles/variants.sh).  Usage: python profiles/build_variants.py name=DEF1,DEF2 name2=DEF ..."""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splatting_b200 import build  # noqa: E402

out_dir = os.path.join(ROOT, "gaussian_splatting_b200", "variants")
os.makedirs(out_dir, exist_ok=True)
for f in os.listdir(out_dir):
    os.remove(os.path.join(out_dir, f))


def one(spec):
    name, defs = spec.split("=", 1)
    return build.build(out=os.path.join(out_dir, name + ".so"), defines=tuple(d for d in defs.split(",") if d))


with ThreadPoolExecutor(4) as ex:
    for r in ex.map(one, sys.argv[1:]):
        print(r)
