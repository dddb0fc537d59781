"""The C-ABI library loads and exports every symbol include/xetgpu.h declares (no GPU needed)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def header_functions():
    txt = open(os.path.join(ROOT, "include", "xetgpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(xetgpu_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_binding_agree():
    from ome_b200 import _lib
    assert header_functions() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    from ome_b200 import _lib
    lib = _lib.load()
    for name in header_functions():
        assert hasattr(lib, name), name


def test_hex_helpers_without_gpu():
    import ome_b200
    h = bytes(range(32))
    s = ome_b200.hash_to_hex(h)
    assert s == "07060504030201000f0e0d0c0b0a090817161514131211101f1e1d1c1b1a1918"
    assert ome_b200.hash_from_hex(s) == h
    with pytest.raises(ome_b200.XetGpuError) as ei:
        ome_b200.hash_from_hex("abc")
    assert ei.value.code == 1  # InvalidConfig, as the reference uses for malformed arguments


def test_no_cpu_fallback():
    """Without a CUDA device the product path must fail loudly, never compute on the CPU."""
    import torch
    import ome_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(ome_b200.XetGpuError) as ei:
        ome_b200.XetHasher()
    assert "no CUDA device" in ei.value.message


def test_product_code_never_touches_oracle():
    pkg = os.path.join(ROOT, "ome_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".go")):
                txt = open(os.path.join(dp, fn), errors="replace").read()
                assert "xet_oracle" not in txt and "oracle/" not in txt, os.path.join(dp, fn)


def test_cpp_consumers_fail_loudly_without_gpu(tmp_path):
    """tools/xetsum (plain C ABI) and tools/xetgpu_host_test (include/xetgpu.hpp) build, link against the
    library and, with no CUDA device, stop with the library's error instead of computing anything."""
    import subprocess
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    p = tmp_path / "x.bin"
    p.write_bytes(b"hello world")
    for exe, args in (("xetsum", [str(p)]), ("xetgpu_host_test", [str(tmp_path)])):
        path = os.path.join(ROOT, "tools", exe)
        if not os.path.exists(path):
            pytest.skip(f"tools/{exe} not built")
        r = subprocess.run([path] + args, capture_output=True, text=True)
        assert r.returncode == 2 and "no CUDA device" in r.stderr, (exe, r.stdout, r.stderr)
