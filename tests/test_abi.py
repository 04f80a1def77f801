"""The C-ABI library loads and exports every symbol include/amber_cuda.h declares, and the
ctypes structs have the layout the C compiler gives them.  No compute calls: runs on CPU."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

from conftest import ROOT


def header_symbols():
    text = open(os.path.join(ROOT, "include", "amber_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(amber_cuda_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from amber_b200 import capi
    assert os.path.exists(capi.LIB_PATH), "libamber_cuda.so is not built (run __graft_entry__.build())"
    lib = ctypes.CDLL(capi.LIB_PATH)
    declared = header_symbols()
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/amber_cuda.h but not exported"
    assert sorted(capi.SYMBOLS) == declared
    assert lib.amber_cuda_abi_version() == capi.ABI_VERSION


def test_struct_layout_matches_the_c_compiler(tmp_path):
    from amber_b200 import capi
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "amber_cuda.h"\nint main(void){\n'
                   'printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(amber_cuda_primitive), sizeof(amber_cuda_material),'
                   'sizeof(amber_cuda_lens), sizeof(amber_cuda_scene_desc), sizeof(amber_cuda_render_params),'
                   'sizeof(amber_cuda_stats), offsetof(amber_cuda_scene_desc, lens), offsetof(amber_cuda_stats, accel));'
                   'return 0;}\n')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [ctypes.sizeof(capi.Primitive), ctypes.sizeof(capi.Material), ctypes.sizeof(capi.Lens),
            ctypes.sizeof(capi.SceneDesc), ctypes.sizeof(capi.RenderParams), ctypes.sizeof(capi.Stats),
            capi.SceneDesc.lens.offset, capi.Stats.accel.offset]
    assert got == want


def test_create_fails_loudly_without_a_gpu():
    """No CPU fallback: on a box without CUDA devices create() must fail with ENODEV."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from amber_b200 import capi
    from amber_b200.render import Device
    with pytest.raises(capi.AmberCudaError, match="no CUDA device"):
        Device(0)


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "amber_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in text and "amber_oracle" not in text and "libamber_ref" not in text, f
