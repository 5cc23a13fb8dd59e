"""CPU: the C-ABI library builds, loads, and exports every symbol include/tutor_b200.h declares.
No compute is called here (there is no GPU in the build container)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "tutor_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tutor_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tutor_b200.h but not exported"


def test_binding_table_covers_header(pkg):
    assert sorted(pkg._lib.SIGNATURES) == declared_symbols()


def test_version_and_error_text(lib):
    assert lib.tutor_version() >= 100
    assert isinstance(lib.tutor_last_error(), bytes)


def test_library_is_sm100a_and_uses_tcgen05(pkg):
    """The product GEMM must be the tcgen05/TMA kernel: look for its SASS mnemonics in the built library."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", pkg._lib.lib_path()], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
        assert mnemonic in sass, mnemonic


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu(pkg, lib):
    cfg = pkg._lib.Gpt2Config(2, 2, 128, 1000, 256, 1e-5, 0)
    h = ctypes.c_void_p()
    rc = lib.tutor_gpt2_create(ctypes.byref(cfg), 0, ctypes.byref(h))
    assert rc == -2 and b"no CPU fallback" in lib.tutor_last_error()
    with pytest.raises(pkg._lib.TutorError):
        pkg.cosine_similarity(torch.ones(1, 8), torch.ones(1, 8))
