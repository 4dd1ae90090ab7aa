"""The JVM-side binding that cannot run here (no JDK, no Scala) is still held to the C ABI: the JNI shim is
type-checked against include/hsgpu.h with a stand-in jni.h, and every native method the Scala object declares
has its JNI function (and the other way round)."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_jni_shim_type_checks_against_the_header():
    r = subprocess.run(["gcc", "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "tests", "stubs"),
                        "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "bindings", "jni", "hsgpu_jni.c")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_every_native_method_has_its_jni_function():
    c = open(os.path.join(ROOT, "bindings", "jni", "hsgpu_jni.c")).read()
    scala = open(os.path.join(ROOT, "bindings", "scala", "com", "microsoft", "hyperspace", "gpu", "HsGpu.scala")).read()
    jni = set(re.findall(r"HS_JNI\(\w+, (\w+)\)", c)) - {"name"}   # "name" is the macro's own parameter
    native = set(re.findall(r"@native def (\w+)\(", scala))
    assert jni == native, (sorted(jni - native), sorted(native - jni))
    # the shim reaches every data-path export of the header
    for name in ("hs_build_index", "hs_build_index_dist", "hs_merge_runs", "hs_join_bucketed", "hs_filter_scan",
                 "hs_hash_bucket", "hs_table_wait", "hs_table_free", "hs_comm_init"):
        assert name + "(" in c, name


def test_jni_makefile_is_a_no_op_without_a_jdk():
    env = dict(os.environ)
    env.pop("JAVA_HOME", None)
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "bindings")], capture_output=True, text=True, env=env)
    assert r.returncode == 0 and "JNI shim not built" in r.stdout
