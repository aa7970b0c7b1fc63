"""The C-ABI library loads and exports every symbol include/hivemall_b200.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

import pytest

import hivemall_b200 as hm
from hivemall_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "hivemall_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hm_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_something():
    names = declared_functions()
    assert "hm_udtf_create" in names and "hm_mhash" in names and len(names) >= 25


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_binding_table_matches_header():
    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_struct_layouts_match_c(tmp_path):
    """sizeof / offsetof as the C compiler sees the header vs the ctypes mirrors."""
    import shutil
    import subprocess
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("cc")
    if not cc:
        pytest.skip("no C compiler")
    fields_c = [n for n, _ in _lib.HmConfig._fields_]
    fields_s = [n for n, _ in _lib.HmStats._fields_]
    src = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void){',
           'printf("%zu\\n", sizeof(hm_config));']
    src += [f'printf("%zu\\n", offsetof(hm_config, {n.rstrip("_")}));' for n in fields_c]
    src += ['printf("%zu\\n", sizeof(hm_stats));']
    src += [f'printf("%zu\\n", offsetof(hm_stats, {n}));' for n in fields_s]
    src += ['return 0;}']
    c = tmp_path / "probe.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "probe"
    subprocess.check_call([cc, str(c), "-o", str(exe)])
    nums = [int(x) for x in subprocess.check_output([str(exe)], text=True).split()]
    want = [ctypes.sizeof(_lib.HmConfig)] + [getattr(_lib.HmConfig, n).offset for n in fields_c]
    want += [ctypes.sizeof(_lib.HmStats)] + [getattr(_lib.HmStats, n).offset for n in fields_s]
    assert nums == want


def test_abi_version_and_error_text():
    lib = _lib.load()
    assert lib.hm_abi_version() == 1
    assert isinstance(lib.hm_last_error(), bytes)
    assert lib.hm_device_count() >= 0


def test_no_cpu_fallback_without_device():
    """Without a CUDA device every compute entry point must fail loudly."""
    if _lib.load().hm_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(hm.HivemallError) as e:
        hm.train_classifier("-dims 16")
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)
    with pytest.raises(hm.HivemallError):
        hm.mhash(["a"])


def test_product_never_imports_the_oracle():
    """Nothing under hivemall_b200/ includes, links, loads or imports the oracle
    (build.py only compiles it for the tests)."""
    pkg = os.path.join(ROOT, "hivemall_b200")
    patterns = [r'#\s*include\s*[<"][^>"]*hm_oracle', r"\bimport\s+_oracle\b", r"\bfrom\s+_oracle\b",
                r"libhm_oracle", r"\bhmo_[a-z_0-9]+\s*\("]
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")) or f == "build.py":
                continue
            text = open(os.path.join(dirpath, f), errors="replace").read()
            for pat in patterns:
                assert not re.search(pat, text), f"{f} matches {pat!r}: the product must not use the oracle"
