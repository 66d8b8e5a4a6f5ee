"""The C-ABI library loads on a machine without a GPU, exports every symbol include/glprover.h declares, and
refuses -- loudly -- to do anything without a CUDA device (there is no CPU path).  CPU only."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from xlayer_prover_b200 import build
    build.build()                                  # nvcc cross-compiles for sm_100a without a GPU
    from xlayer_prover_b200._capi import load
    return load()


def declared_functions():
    text = open(os.path.join(ROOT, "include", "glprover.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(glp_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    from xlayer_prover_b200._capi import SIGNATURES
    assert declared_functions() == sorted(SIGNATURES)


def test_every_declared_symbol_is_exported(lib):
    for name in declared_functions():
        assert hasattr(lib.cdll, name), name


def test_host_only_entry_points(lib):
    assert lib.glp_tree_num_elements(64) == (2 * 64 - 1) * 4
    assert lib.glp_tree_num_elements(0) == 0
    assert lib.glp_root(6, 1) == 8                  # W[6] = 8
    assert lib.glp_root(1, 1) == 0xFFFFFFFF00000000  # W[1] = -1
    assert lib.glp_root(32, 1) == 7277203076849721926
    assert lib.glp_root(4, 16) == 1


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from xlayer_prover_b200._capi import GLP_E_CUDA
    h = ctypes.c_void_p()
    rc = lib.glp_init(0, ctypes.byref(h))
    assert rc == GLP_E_CUDA and not h.value
    assert b"no CPU path" in lib.glp_last_error()
    from xlayer_prover_b200 import Context, GlpError
    with pytest.raises(GlpError):
        Context(0)


def test_product_does_not_import_oracle():
    """Nothing under the package may reference oracle/ or the emulator."""
    pkg = os.path.join(ROOT, "xlayer_prover_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "refbind" not in text and "gl_oracle" not in text and "libglref" not in text, f
                if f.endswith(".py"):
                    assert "libglprover_emu" not in text, f
