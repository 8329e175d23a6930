"""The C-ABI shared library: builds for sm_100a, loads, exports every symbol include/svtk.h declares, and
refuses to compute without a GPU (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "svtk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(svtk_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_the_seams():
    syms = declared_symbols()
    for s in ("svtk_deletion", "svtk_insertion", "svtk_inversion", "svtk_refine_windows", "svtk_process_vcf",
              "svtk_consensus_pos_batch", "svtk_sliding_window_ins", "svtk_load_reads"):
        assert s in syms


def test_library_exports_every_declared_symbol(lib_built):
    lib = ctypes.CDLL(lib_built)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/svtk.h but not exported"


def test_python_binding_lists_every_symbol():
    from svtrek_b200 import api
    assert sorted(api.EXPORTS) == declared_symbols()


def test_sass_is_sm100a(lib_built):
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", lib_built], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_no_gpu_fails_loudly(lib_built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from svtrek_b200 import api
    with pytest.raises(api.SvtkError) as e:
        api.Auditor(0)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_product_does_not_touch_the_oracle():
    """Nothing under svtrek_b200/ may import, link or name the oracle."""
    pkg = os.path.join(ROOT, "svtrek_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower().replace("no oracle", ""), os.path.join(dirpath, f)
