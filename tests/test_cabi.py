"""CPU: the C-ABI shared library builds, loads and exports exactly what include/gd_b200.h declares.
No compute is launched here (no GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from gesturediffusion_b200 import build
    build.build()
    from gesturediffusion_b200 import _cabi
    return _cabi.load()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gd_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    from gesturediffusion_b200 import _cabi
    declared = _declared_symbols()
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), "libgd_b200.so does not export %s" % name
    assert sorted(_cabi.EXPORTED_SYMBOLS) == declared


def test_abi_version_and_error_string(lib):
    from gesturediffusion_b200 import _cabi
    assert lib.gd_abi_version() == 1
    ep = _cabi.GdEpilogue()
    rc = lib.gd_gemm_f32(None, 8, None, 8, 4, 4, 8, ctypes.byref(ep), None)
    assert rc == -1                                     # GD_ERR_INVALID, nothing launched
    assert b"null pointer" in lib.gd_last_error()
    assert lib.gd_launch_count() == 0


def test_struct_layout_matches_header(lib, tmp_path):
    """The ctypes mirrors must have exactly the C layout: compile the header with gcc and compare
    sizeof / offsetof of every field."""
    import shutil
    import subprocess
    from gesturediffusion_b200 import _cabi
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "gd_b200.h"', 'int main(void){']
    for cname, cls in (("gdEpilogue", _cabi.GdEpilogue), ("gdPosteriorArgs", _cabi.GdPosteriorArgs)):
        lines.append('printf("%s %%zu\\n", sizeof(%s));' % (cname, cname))
        for fname, _ in cls._fields_:
            lines.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (cname, fname, cname, fname))
    lines.append('return 0;}')
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True).stdout.splitlines())
    for cname, cls in (("gdEpilogue", _cabi.GdEpilogue), ("gdPosteriorArgs", _cabi.GdPosteriorArgs)):
        assert int(got[cname]) == ctypes.sizeof(cls)
        for fname, _ in cls._fields_:
            assert int(got["%s.%s" % (cname, fname)]) == getattr(cls, fname).offset, (cname, fname)


def test_no_cpu_fallback():
    import torch
    from gesturediffusion_b200 import _cabi, kernels as K
    with pytest.raises(_cabi.GdError):
        K.layernorm(torch.zeros(4, 128), torch.ones(128), torch.zeros(128), out_f32=torch.zeros(4, 128))


def test_library_sass_is_blackwell_native():
    """The GEMM must be tcgen05/TMA code: UTC*MMA, LDTM and UTMALDG in the sm_100a SASS."""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not available")
    from gesturediffusion_b200 import _cabi
    sass = subprocess.run(["cuobjdump", "-sass", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "LDTM", "UTMALDG"):
        assert mnemonic in sass, mnemonic
