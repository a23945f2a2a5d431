"""Kernel experiments: build flybody_b200/_exp/libflyb_<name>.so with extra -D flags (all kernel configurations, compiled in parallel).
   usage: python tools/build_variant.py <name> [-DFLAG ...]   ; run with FLYB_LIB_PATH=<that file>"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flybody_b200 import engine as E

name, flags = sys.argv[1], sys.argv[2:]
out = os.path.join(os.path.dirname(E.LIB_PATH), "_exp", f"libflyb_{name}.so")
os.makedirs(os.path.dirname(out), exist_ok=True)
print(E.compile_library(out, extra_flags=flags))
