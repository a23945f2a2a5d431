"""The C-ABI boundary without a GPU: header generation is current, struct layouts agree, every declared symbol is exported."""
import ctypes
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_model_header_is_current():
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_header
    assert gen_header.render() == open(os.path.join(ROOT, "include", "flyb_model.h")).read()


def test_struct_size_matches_c(tmp_path):
    from flybody_b200.model_pack import FlybModel
    src = tmp_path / "s.c"
    src.write_text('#include <stdio.h>\n#include "flyb_model.h"\nint main(){printf("%zu", sizeof(FlybModel));return 0;}\n')
    exe = tmp_path / "s"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    assert int(subprocess.check_output([str(exe)])) == ctypes.sizeof(FlybModel)


def test_library_exports_every_declared_symbol():
    from flybody_b200 import engine
    engine.build_library()
    hdr = open(os.path.join(ROOT, "include", "flyb.h")).read()
    declared = sorted(set(re.findall(r"\b(flyb_[a-z_]+)\s*\(", hdr)))
    assert declared == sorted(engine.EXPORTS)
    L = ctypes.CDLL(engine.LIB_PATH)
    for name in declared:
        assert getattr(L, name) is not None


def test_create_fails_loudly_without_a_gpu(models):
    import torch
    if torch.cuda.is_available():
        return
    from flybody_b200.engine import Engine
    try:
        Engine(models["walk_on_ball"], 4)
    except RuntimeError as e:
        assert "no CUDA device" in str(e) or "CUDA" in str(e)
    else:
        raise AssertionError("engine creation must fail without a GPU (no CPU fallback)")


def test_product_package_does_not_touch_the_oracle():
    """oracle/ and tests/emu are checkers: nothing under flybody_b200/ may import, include, link or load them."""
    pkg = os.path.join(ROOT, "flybody_b200")
    bad = re.compile(r"(^|\n)\s*(import|from)\s+(oracle|tests)\b|#include[^\n]*(oracle|emu)|libflyb_oracle|libflyb_emu|oracle/_")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert not bad.search(open(os.path.join(dp, f)).read()), f
