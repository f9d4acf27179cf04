"""CPU: the C-ABI library loads without a GPU, exports every symbol include/psg.h declares, and fails loudly (no CPU
fallback) when asked to compute without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import psgradientsdf_b200 as psg


def declared_symbols():
    text = open(psg.HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(psg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = psg.lib()
    names = declared_symbols()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, f"declared in include/psg.h but not exported: {missing}"


def test_header_cites_the_reference():
    text = open(psg.HEADER_PATH).read()
    for needle in ("VolumetricGradSdf.cpp:51-138", "Optimizer.cpp:30-47", "OptimizerAux.cpp:237-257", "PsOptimizer.cpp:85-234",
                   "main_ps.cpp:323-326"):
        assert needle in text


def test_argument_validation_needs_no_gpu():
    L = psg.lib()
    h = C.c_void_p()
    dims = np.array([0, 4, 4], np.int32)
    st = L.psg_volume_create(C.byref(h), dims, 0.01, np.zeros(3, np.float32), 0.05, 0.5, 3.5, 64, 0)
    assert st == 1 and b"not supported" in L.psg_last_error()
    assert L.psg_profile_kernels() > 10 and L.psg_profile_name(2) == b"k_integrate"
    assert L.psg_kernel_launches() >= 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    v = psg.VolumetricGradSdf([8, 8, 8], 0.01, [0, 0, 0], 0.05)
    with pytest.raises(psg.PsgError, match="cuda|CUDA|device"):
        v.download()


def test_product_does_not_import_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "psgradientsdf_b200")
    offenders = []
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, fn), errors="replace").read()
                if re.search(r"(from|import)\s+oracle|oracle/|psg_oracle|libpsgref|pyref|pyoracle", src):
                    offenders.append(os.path.join(dirpath, fn))
    assert not offenders, f"product sources reference the oracle: {offenders}"


def test_header_is_plain_c(tmp_path):
    """include/psg.h is the drop-in boundary for a C / C++ host: it must compile as C99 without any CUDA or C++ header, and
    a C program must link against the library."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        import pytest
        pytest.skip("no gcc")
    src = tmp_path / "t.c"
    src.write_text('#include "psg.h"\n#include <stdio.h>\n'
                   'int main(void) { psg_iteration_config c; psg_settings s; int n = -1; (void)c; (void)s;\n'
                   '  if (psg_shard_blocks(100, 1, 2, &n, &n, &n) != 0) return 2;\n'
                   '  if (psg_volume_create(0, 0, 0.f, 0, 0.f, 0.f, 0.f, 1, 0) == 0) return 3;   /* null arguments are refused */\n'
                   '  printf("%s\\n", psg_last_error()); return 0; }\n')
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    inc = os.path.join(ROOT, "include")
    subprocess.check_call([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-fsyntax-only", str(src)])
    libdir = os.path.join(ROOT, "psgradientsdf_b200")
    exe = tmp_path / "t"
    subprocess.check_call([gcc, "-std=c99", "-I", inc, str(src), "-o", str(exe), "-L", libdir, "-lpsg_b200", f"-Wl,-rpath,{libdir}"])
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "null" in r.stdout
