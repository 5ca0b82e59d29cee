"""The C-ABI library loads on a machine without a GPU and exports every symbol include/smi.h declares
(no compute calls here)."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "smi.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^(?:int|size_t|long long|const char\*)\s+(smi_\w+)\s*\(", src, flags=re.M)
    assert len(names) >= 25
    return names


def test_library_exports_every_declared_symbol(lib):
    from smi_experiments_b200 import _lib
    for name in _declared():
        assert hasattr(lib, name), "libsmi.so does not export %s" % name
        assert name in _lib.PROTOTYPES, "_lib.PROTOTYPES does not bind %s" % name
    assert set(_lib.PROTOTYPES) == set(_declared())


def test_version_and_error_channel(lib):
    assert lib.smi_version() == 100
    assert isinstance(lib.smi_last_error(), bytes)
    # argument validation happens before any CUDA call: safe without a GPU
    assert lib.smi_stein_force_f32(None, None, 4, 4, 1.0, 1.0, 0, 0.0, 0, None, None, None, None, 0, None) == -1
    assert b"null pointer" in lib.smi_last_error()
    assert lib.smi_stein_force_workspace_bytes(20, 6604, 0) > 0
    assert lib.smi_stein_force_workspace_bytes(16384, 10000, 0) > 4 * 16384 * 16384
    assert lib.smi_launch_count(1) >= 0


def test_model_desc_struct_matches_header():
    from smi_experiments_b200 import _lib
    # 11 int32/float fields + 16 sites x 5 int32 (include/smi.h SmiModelDesc)
    assert C.sizeof(_lib.SmiSite) == 20
    assert C.sizeof(_lib.SmiModelDesc) == 11 * 4 + 16 * 20


def test_no_oracle_in_product_package():
    """The product path never imports the CPU oracle (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "smi_experiments_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f
