"""CPU: the C-ABI library loads and exports every symbol include/gmr_b200.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "gmr_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gmr_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from gmr_b200 import _lib
    from gmr_b200.build import build
    build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), f"{n} declared in the header but not exported"
    assert sorted(_lib.EXPORTS) == names
    assert L.gmr_abi_version() == 2


def test_struct_layout_matches_header():
    """sizeof(gmr_model_desc) computed by ctypes equals the C compiler's."""
    import subprocess
    import tempfile
    from gmr_b200.abi import gmr_diag, gmr_model_desc
    src = '#include <stdio.h>\n#include "gmr_b200.h"\nint main(){printf("%zu %zu %zu %zu\\n", sizeof(gmr_model_desc), sizeof(gmr_diag), ' \
          '__builtin_offsetof(gmr_model_desc, damping), __builtin_offsetof(gmr_model_desc, task_wrot));' \
          'printf("%zu %zu\\n", __builtin_offsetof(gmr_model_desc, dof_vmax), __builtin_offsetof(gmr_diag, status_dev));return 0;}\n'
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(td, "t")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        a, b, off1, off2, off3, off4 = map(int, subprocess.check_output([exe]).split())
    assert ctypes.sizeof(gmr_model_desc) == a
    assert ctypes.sizeof(gmr_diag) == b
    assert gmr_model_desc.damping.offset == off1
    assert gmr_model_desc.task_wrot.offset == off2
    assert gmr_model_desc.dof_vmax.offset == off3
    assert gmr_diag.status_dev.offset == off4


def test_compile_time_layouts_are_current():
    """csrc/gmr_layouts.cuh (generated) holds exactly the layouts the library's make_layout computes for the
    constructible pairs; a stale file would silently send those robots to the generic kernel."""
    import subprocess
    import sys
    from gmr_b200.build import build
    build()
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_layouts.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def test_engine_fails_loudly_without_library(monkeypatch, tmp_path):
    import pytest
    from gmr_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "missing.so"))
    with pytest.raises(_lib.EngineError):
        _lib.lib()
