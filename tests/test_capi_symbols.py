"""-m "not gpu": the C-ABI library builds for sm_100a, loads without a GPU, exports every symbol that
include/ink_capi.h declares, and refuses (loudly) to create a context when no CUDA device is present —
there is no CPU fallback in the product."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ink_capi.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ink_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "ink_simulation_b200", "csrc")], stdout=subprocess.DEVNULL)
    from ink_simulation_b200 import _capi

    return _capi.load()


def test_header_symbols_are_exported(lib):
    from ink_simulation_b200 import _capi

    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ink_capi.h but not exported"
    # and the ctypes table covers exactly the header
    assert sorted(_capi.SYMBOLS) == names


def test_struct_layout_matches_header(lib):
    """sizeof(ink_params)/sizeof(ink_stats) as laid out by ctypes must match what the C compiler sees."""
    from ink_simulation_b200 import _capi

    code = '#include "ink_capi.h"\n#include <stdio.h>\nint main(){printf("%zu %zu\\n", sizeof(ink_params), sizeof(ink_stats));return 0;}'
    exe = os.path.join(ROOT, "tests", "_sizeof_probe")
    try:
        subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=code.encode(), check=True)
        a, b = map(int, subprocess.check_output([exe]).split())
    finally:
        if os.path.exists(exe):
            os.unlink(exe)
    assert ctypes.sizeof(_capi.InkParams) == a
    assert ctypes.sizeof(_capi.InkStats) == b


def test_default_params_are_the_reference_globals(lib):
    from ink_simulation_b200 import _capi

    p = _capi.default_params()
    assert (p.X, p.Y, p.Z) == (20, 45, 20)                      # src/system.h:15-17
    assert p.cell_dim == 1.0 and p.buffer_size == 3              # :22-23
    assert p.density == np.float32(0.95) and p.k_vort == 1.0     # :25,27
    assert p.viscosity == np.float32(1.0016) and p.p_atm == 1.0  # :30-31
    assert tuple(p.gravity) == (0.0, np.float32(-0.1), 0.0)      # :34
    assert p.k_cfl == np.float32(0.2) and p.dt_min == np.float32(0.005) and p.dt_max == 1.0  # :36-38


def test_no_cpu_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from ink_simulation_b200 import InkError, System

    with pytest.raises(InkError, match="no CUDA device"):
        System(8, 25, 8)


def test_product_does_not_import_oracle():
    """The product path must never route through the checker."""
    for base, _, files in os.walk(os.path.join(ROOT, "ink_simulation_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(base, f)).read()
                assert "import oracle" not in txt and "libinkoracle" not in txt and "libinkref" not in txt, f


def test_every_entry_point_is_mapped_to_the_reference_in_the_docs():
    """INTEGRATION.md is what a maintainer of the reference binds against: every entry point of the header must appear there
    (with the reference member it replaces, or 'no reference equivalent'), and the header itself must cite reference lines."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = [n for n in declared_symbols() if n not in doc]
    assert not missing, missing
    assert len(re.findall(r"src/[A-Za-z]+\.(?:cpp|h):\d+", open(HEADER).read())) >= 15
