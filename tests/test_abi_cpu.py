"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every
symbol include/muscle_b200.h declares (and the ctypes table mirrors the header one to one); the
size helpers answer without a GPU; the product package never touches oracle/ and fails loudly
instead of falling back to the CPU."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "muscle_b200.h")


@pytest.fixture(scope="module")
def lib():
    from muscle_b200 import build, _lib
    build.build()
    return _lib.load()


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", "", src, flags=re.M)
    names = re.findall(r"\b(muscle_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_and_ctypes_table_agree(lib):
    from muscle_b200 import _lib
    declared = _declared_functions()
    assert declared, "no functions parsed from the header"
    assert declared == _lib.exported_symbols()
    for name in declared:
        assert getattr(lib, name) is not None


def test_library_exports_are_plain_c(lib):
    from muscle_b200 import _lib
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    for name in _declared_functions():
        assert name in exported, name
    # nothing torch-typed crosses the boundary: the library links against the CUDA runtime only
    ldd = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in ldd and "c10" not in ldd


def test_diag_fixture_is_a_library_of_its_own(lib):
    """muscle_debug_tile (include/muscle_b200_diag.h) is a test fixture: built, but not inside the product library"""
    from muscle_b200 import _lib
    diag = _lib.load_diag()
    assert diag.muscle_debug_tile is not None
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "muscle_debug_tile" not in out and "debug_tile_kernel" not in out


def test_size_helpers_without_gpu(lib):
    assert lib.muscle_version() >= 200
    assert lib.muscle_pad_rows(1) == 256 and lib.muscle_pad_rows(256) == 256 and lib.muscle_pad_rows(257) == 512
    assert lib.muscle_pad_rows(0) == 256
    assert lib.muscle_infonce_part_slots() >= 2
    small = lib.muscle_dense_infonce_workspace_bytes(2048, 128, 21)
    big = lib.muscle_dense_infonce_workspace_bytes(25088, 128, 21)
    assert 0 < small < big
    assert lib.muscle_dense_infonce_workspace_bytes(2048, 96, 21) == 0       # unsupported dim
    assert lib.muscle_dense_infonce_workspace_bytes(2048, 128, 99) == 0      # too many classes
    assert lib.muscle_infonce_prepare_scratch_bytes(25088) > 0


def test_no_cpu_fallback():
    import muscle_b200 as M
    with pytest.raises(RuntimeError):
        M.dense_infonce(torch.randn(64, 128), torch.zeros(64, dtype=torch.int64))
    if not torch.cuda.is_available():
        from muscle_b200.stages import CudaStages
        with pytest.raises(RuntimeError):
            CudaStages("cpu")


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "muscle_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), os.path.join(dirpath, f)
                assert "/root/reference" not in text, os.path.join(dirpath, f)
    code = "import sys; import muscle_b200; assert not any(m.split('.')[0]=='oracle' for m in sys.modules)"
    subprocess.run([sys.executable, "-c", code], cwd=ROOT, check=True)
