"""The JNI shim a maintainer builds next to libfsx (hops-examples_b200/jvm/fsx_jni.c) cannot be linked into a JVM in this
image (no JDK), but it can be type-checked: it compiles against a minimal stand-in jni.h with -Wall -Werror, links against
libfsx.so, and exports one function per @native method of jvm/Native.scala."""
import os
import re
import subprocess

from conftest import ROOT

JVM = os.path.join(ROOT, "hops-examples_b200", "jvm")


def test_jni_shim_compiles_and_covers_every_native_method(tmp_path):
    so = tmp_path / "libfsx_jni.so"
    cmd = ["gcc", "-shared", "-fPIC", "-std=c11", "-Wall", "-Werror", "-Wno-unused-parameter", "-I", os.path.join(ROOT, "tests", "mock_jni"),
           "-I", os.path.join(ROOT, "include"), os.path.join(JVM, "fsx_jni.c"), "-L", os.path.join(ROOT, "hops-examples_b200", "lib"), "-lfsx", "-o", str(so)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    exported = subprocess.run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True).stdout
    scala = open(os.path.join(JVM, "Native.scala")).read()
    methods = re.findall(r"@native def (\w+)\(", scala)
    assert len(methods) >= 30
    for m in methods:
        assert f"Native_00024_{m}\n" in exported or f"Native_00024_{m}" in exported.split(), f"fsx_jni.c lacks the function for @native def {m}"
