"""jni/ftf_jni.cc cannot be built without a JDK (none in this image), but it must not rot: it is type-checked here against
tests/jni_check/jni.h, a declaration-only subset of the JNI header, and checked to bind every native method that
jvm/.../Native.java declares -- and nothing else."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shim_type_checks_against_the_c_abi():
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "tests", "jni_check"),
           "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "jni", "ftf_jni.cc")]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout
    # and without any jni.h the file is an empty translation unit, not an error
    p = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "jni", "ftf_jni.cc")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout


def test_every_native_method_has_exactly_one_binding():
    java = open(os.path.join(ROOT, "jvm", "src", "main", "java", "org", "apache", "flink", "contrib", "tensorflow", "b200", "Native.java")).read()
    declared = set(re.findall(r"public static native [\w\[\]]+\s+(\w+)\(", java))
    shim = open(os.path.join(ROOT, "jni", "ftf_jni.cc")).read()
    bound = set(re.findall(r"NATIVE\(\w+, (\w+)\)", shim)) - {"name"}      # the macro's own definition
    assert declared and declared == bound, (sorted(declared - bound), sorted(bound - declared))
    # each binding reaches the C ABI: every ftf_* name the shim mentions is declared in include/ftf.h
    header = open(os.path.join(ROOT, "include", "ftf.h")).read()
    for name in set(re.findall(r"\b(ftf_[a-z0-9_]+)\(", shim)):
        assert re.search(r"\b%s\(" % name, header), name
