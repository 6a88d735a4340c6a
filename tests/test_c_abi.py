"""The public header is plain C99 (and C++), and the C example links against the built library."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")


@pytest.mark.parametrize("compiler,flags", [("gcc", ["-std=c99", "-pedantic", "-x", "c"]),
                                            ("g++", ["-std=c++17", "-x", "c++"])])
def test_header_compiles_as_c_and_cxx(tmp_path, compiler, flags):
    if not shutil.which(compiler):
        pytest.skip(f"{compiler} not installed")
    src = tmp_path / "t.c"
    src.write_text('#include "gadmm_b200.h"\n'
                   "int main(void) { GadmmChain c; GadmmStar s; GadmmLogistic l; (void)c; (void)s; (void)l;\n"
                   "  return gadmm_packed_elems(64) > 0 ? 0 : 1; }\n")
    r = subprocess.run([compiler] + flags + ["-Wall", "-Wextra", "-Werror", "-fsyntax-only", "-I", INC, str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_c_example_links_against_the_library(tmp_path):
    cuda = "/usr/local/cuda"
    if not (shutil.which("gcc") and os.path.exists(os.path.join(cuda, "include", "cuda_runtime_api.h"))):
        pytest.skip("gcc or the CUDA toolkit headers are missing")
    from gadmm_b200._native import build, lib_path
    build()
    libdir = os.path.dirname(lib_path())
    out = tmp_path / "demo"
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-I", INC, "-I", os.path.join(cuda, "include"),
                        os.path.join(ROOT, "examples", "c_api_demo.c"), "-L", libdir, "-lgadmm_b200",
                        "-L", os.path.join(cuda, "lib64"), "-lcudart", f"-Wl,-rpath,{libdir}", "-o", str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert out.exists()
