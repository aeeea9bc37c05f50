"""The JNI shim (java/langstream-cuda/src/main/c) cannot be built here - no JDK in this image - but it can be type-checked
against the C ABI header: tests/jni_stub/jni.h declares the handful of JNI functions the shim uses, so a call that no
longer matches include/langstream_cuda.h (an argument added, a type changed) fails this test instead of a maintainer's build."""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_jni_shim_type_checks_against_the_header():
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    shim = ROOT / "java" / "langstream-cuda" / "src" / "main" / "c" / "langstream_cuda_jni.c"
    r = subprocess.run([gcc, "-fsyntax-only", "-Wall", "-Werror", "-I", str(ROOT / "tests" / "jni_stub"), "-I", str(ROOT / "include"),
                        str(shim)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_every_native_method_has_a_shim_function():
    """`public static native` methods of LangstreamCuda.java and the Java_..._<name> functions of the shim, one to one."""
    import re
    java = (ROOT / "java" / "langstream-cuda" / "src" / "main" / "java" / "ai" / "langstream" / "cuda" / "LangstreamCuda.java").read_text()
    shim = (ROOT / "java" / "langstream-cuda" / "src" / "main" / "c" / "langstream_cuda_jni.c").read_text()
    natives = set(re.findall(r"public static native [\w.<>\[\]]+\s+(\w+)\s*\(", java))
    funcs = set(re.findall(r"Java_ai_langstream_cuda_LangstreamCuda_(\w+)\s*\(", shim))
    assert natives == funcs, (sorted(natives - funcs), sorted(funcs - natives))
