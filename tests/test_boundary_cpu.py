"""The reference-side binding (integration/): the JNI shim compiles (against tests/jni_stub/jni.h — the image has no JDK), wraps
every pg_* entry point of include/prestogpu.h, and agrees name for name with the `native` declarations of PrestoGpu.java and with
the natives the Java operator sources call.  No GPU, no compute calls."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
JNI = os.path.join(ROOT, "integration", "jni")
JAVA = os.path.join(ROOT, "integration", "java", "io", "trino", "gpu")
HEADER = os.path.join(ROOT, "include", "prestogpu.h")


def _strip_comments(text):
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def _header_functions():
    return sorted(set(re.findall(r"PG_API\s+[\w\s\*]+?\b(pg_\w+)\s*\(", _strip_comments(open(HEADER).read()))))


@pytest.fixture(scope="module")
def shim_objects(tmp_path_factory):
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("gcc not available")
    out = tmp_path_factory.mktemp("jni")
    objs = []
    for name in ("prestogpu_jni.c", "prestogpu_jni_ops.c"):
        obj = str(out / (name[:-2] + ".o"))
        r = subprocess.run([gcc, "-std=c11", "-Wall", "-Wextra", "-Werror", "-Wno-unused-parameter", "-fPIC", "-c", "-I", os.path.join(ROOT, "tests", "jni_stub"),
                            "-I", os.path.join(ROOT, "include"), os.path.join(JNI, name), "-o", obj], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]
        objs.append(obj)
    return objs


def _symbols(objs, kind):
    r = subprocess.run(["nm"] + objs, capture_output=True, text=True)
    assert r.returncode == 0
    return sorted(set(line.split()[-1] for line in r.stdout.splitlines() if len(line.split()) >= 2 and line.split()[-2] == kind))


def test_shim_compiles_and_links_into_a_shared_object_against_the_library_symbols(shim_objects, tmp_path):
    """Every undefined pg_* symbol of the shim is an entry point the header declares (so linking against libprestogpu.so resolves)."""
    undefined = [s for s in _symbols(shim_objects, "U") if s.startswith("pg_")]
    declared = set(_header_functions())
    assert undefined and not [s for s in undefined if s not in declared]


def test_every_entry_point_has_a_jni_wrapper(shim_objects):
    used = set(s for s in _symbols(shim_objects, "U") if s.startswith("pg_"))
    missing = [f for f in _header_functions() if f not in used]
    assert not missing, "pg_* functions without a JNI wrapper: %s" % missing


def test_java_natives_match_the_exported_jni_symbols(shim_objects):
    exported = set(s[len("Java_io_trino_gpu_PrestoGpu_"):] for s in _symbols(shim_objects, "T") if s.startswith("Java_io_trino_gpu_PrestoGpu_"))
    java = _strip_comments(open(os.path.join(JAVA, "PrestoGpu.java")).read())
    declared = set(re.findall(r"public static native [\w\.\[\]]+ (\w+)\(", java))
    assert exported and exported == declared, (sorted(exported - declared), sorted(declared - exported))


def test_native_parameter_counts_match():
    """The number of parameters of every native declaration equals that of its C definition (after JNIEnv*, jclass)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_jni_natives
    java = _strip_comments(open(os.path.join(JAVA, "PrestoGpu.java")).read())
    for ret, name, plist in gen_jni_natives.natives():
        m = re.search(r"public static native ([\w\.\[\]]+) %s\((.*?)\);" % name, java)
        assert m, name
        assert m.group(1) == ret, name
        assert [p.strip() for p in m.group(2).split(",") if p.strip()] == plist, name


def test_java_sources_only_call_declared_natives_and_classes_exist():
    java = _strip_comments(open(os.path.join(JAVA, "PrestoGpu.java")).read())
    declared = set(re.findall(r"public static native [\w\.\[\]]+ (\w+)\(", java)) | {"check", "lastError"}
    constants = set(re.findall(r"public static final int (\w+)", java))
    want_classes = ["GpuHashAggregationOperator", "GpuHashAggregationOperatorFactory", "GpuHashBuilderOperator", "GpuHashBuilderOperatorFactory",
                    "GpuLookupJoinOperator", "GpuLookupJoinOperatorFactory", "GpuGroupByHash", "GpuGroupedAccumulator", "GpuLookupSource", "BlockFields",
                    "BlockArrays", "PageArrays", "OutputArrays", "PrestoGpu",
                    "GpuMarkDistinctOperator", "GpuDistinctLimitOperator", "GpuSetBuilderOperator", "GpuHashSemiJoinOperator", "GpuLookupOuterOperator"]
    for c in want_classes:
        path = os.path.join(JAVA, c + ".java")
        assert os.path.exists(path), c
        text = _strip_comments(open(path).read())
        assert re.search(r"\b(class|interface)\s+%s\b" % c, text), c
        assert text.count("{") == text.count("}") and text.count("(") == text.count(")"), "unbalanced braces in %s" % c
        for ref in re.findall(r"PrestoGpu\.(\w+)", text):
            assert ref in declared or ref in constants, "%s.java uses PrestoGpu.%s which is not declared" % (c, ref)
