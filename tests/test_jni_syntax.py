"""jni/ytk_jni.cpp is the glue a maintainer compiles inside ytk-learn (there is no JDK in this image): it
must at least parse and type-check against the C ABI.  jni/stub/jni.h declares the JNI names it uses."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_jni_glue_parses_against_the_header():
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wno-comment", "-I", os.path.join(ROOT, "jni", "stub"),
                        "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "jni", "ytk_jni.cpp")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_every_java_native_has_its_jni_function():
    java = open(os.path.join(ROOT, "jni", "GpuTreeMaker.java")).read()
    cpp = open(os.path.join(ROOT, "jni", "ytk_jni.cpp")).read()
    natives = re.findall(r"private static native \w+ (\w+)\(", java)
    assert len(natives) >= 8
    for n in natives:
        assert re.search(r"YTK_JNI\(\w+, %s\)" % n, cpp), n
    assert "null, null" not in java          # the sampled features and rows go down to the library
