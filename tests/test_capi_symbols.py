"""The C-ABI library loads on a GPU-less box and exports every function include/*.h declares
(no compute calls here)."""
import ctypes
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    names = []
    for header in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(header).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_declares_the_documented_entry_points():
    names = _declared_functions()
    for required in ("b200_gemm_bf16", "b200_gemm_f32", "b200_attention_fwd", "b200_ffn_fwd", "b200_last_error",
                     "b200_version"):
        assert required in names


def test_library_exports_every_declared_symbol(ljs):
    lib = ljs.capi.load()
    for name in _declared_functions():
        assert hasattr(lib, name), f"libb200sl.so does not export {name}"
    assert lib.b200_version() == 100
    assert set(ljs.capi.EXPORTED_SYMBOLS) <= set(_declared_functions())


def test_argument_validation_needs_no_gpu(ljs):
    """Bad arguments are rejected before any CUDA call, with a message in b200_last_error()."""
    lib = ljs.capi.load()
    rc = lib.b200_gemm_bf16(None, 8, None, 8, 0, None, None, 8, 0, 8, 8, 8, 0, None)
    assert rc == 1 and b"null operand" in lib.b200_last_error()
    buf = ctypes.create_string_buffer(64)
    p = ctypes.cast(buf, ctypes.c_void_p)
    rc = lib.b200_gemm_bf16(p, 8, p, 8, 0, None, p, 8, 0, 8, 12, 8, 0, None)
    assert rc == 1 and b"multiples of 8" in lib.b200_last_error()
    rc = lib.b200_gemm_bf16(p, 8, p, 8, 0, None, p, 8, 0, 8, 8, 8, 3, None)
    assert rc == 1 and b"needs a bias" in lib.b200_last_error()


def test_missing_library_fails_loudly(ljs, monkeypatch, tmp_path):
    monkeypatch.setattr(ljs.capi, "_lib", None)
    monkeypatch.setattr(ljs.capi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ljs.capi.B200Error, match="no CPU/PyTorch fallback"):
        ljs.capi.load()


def test_mesh_create_rejects_bad_shapes_before_touching_nccl(ljs):
    """Argument checks of b200_mesh_create run on the host before any NCCL call (no GPU needed)."""
    import ctypes
    capi = ljs.capi
    lib = capi.load()
    mesh = ctypes.c_void_p()
    uid = ctypes.create_string_buffer(128)
    for shape, world, rank in [((2, 4), 4, 0), ((2, 0), 0, 0), ((2, 2), 4, 7)]:
        arr = (ctypes.c_int * len(shape))(*shape)
        rc = lib.b200_mesh_create(uid, world, rank, arr, len(shape), ctypes.byref(mesh))
        assert rc == 1, (shape, world, rank)                       # B200_ERR_INVALID_ARGUMENT
        assert b"b200_mesh_create" in lib.b200_last_error()
    assert lib.b200_comm_size(None) == -1 and lib.b200_mesh_destroy(None) == 0
