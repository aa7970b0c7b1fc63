"""The JNI glue of INTEGRATION.md section 3 (bindings/jni/): no JDK exists in this image, so the glue is
(1) type-checked against include/hivemall_b200.h with a stand-in of the JNI declarations it uses,
(2) checked name by name against the native methods of HivemallB200.java, and (3, GPU) RUN without a JVM: a
hand-made JNIEnv drives the Java_* functions through create / hostAlloc / process / close / forward /
destroy and the resulting model table must equal, bit for bit, what the Python binding trains from the same
rows.  What stays unverified is only the JVM's side of the JNI contract.
"""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
JNI = os.path.join(ROOT, "bindings", "jni")
CC = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-fPIC", "-I" + os.path.join(JNI, "compile_check"),
      "-I" + os.path.join(ROOT, "include")]


def test_glue_compiles_and_covers_every_native_method(tmp_path):
    obj = str(tmp_path / "glue.o")
    subprocess.check_call(CC + ["-c", os.path.join(JNI, "hivemall_b200_jni.c"), "-o", obj])
    syms = subprocess.check_output(["nm", obj], text=True)
    defined = set(re.findall(r" T (Java_\w+)", syms))
    java = open(os.path.join(JNI, "HivemallB200.java")).read()
    natives = re.findall(r"public static native \S+ (\w+)\(", java)
    assert len(natives) >= 14
    for name in natives:
        assert f"Java_hivemall_gpu_HivemallB200_{name}" in defined, name
    assert len(defined) == len(natives)
    # every library entry point the glue calls is declared in the header (and so exported: tests/test_abi.py)
    header = open(os.path.join(ROOT, "include", "hivemall_b200.h")).read()
    for fn in set(re.findall(r"\b(hm_\w+)\(", open(os.path.join(JNI, "hivemall_b200_jni.c")).read())):
        assert re.search(r"\b%s\(" % fn, header), fn


@pytest.mark.gpu
def test_glue_runs_without_a_jvm_and_matches_the_python_binding(tmp_path):
    import hivemall_b200 as hm
    exe = str(tmp_path / "jni_driver")
    libdir = os.path.dirname(hm.LIB_PATH)
    subprocess.check_call(CC + [os.path.join(JNI, "hivemall_b200_jni.c"), os.path.join(JNI, "compile_check", "driver.c"),
                                "-L" + libdir, "-lhivemall_b200", "-Wl,-rpath," + libdir, "-o", exe])
    options = "-opt adagrad -mini_batch 64 -dims 1024"
    out = subprocess.run([exe, options], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    lines = out.stdout.splitlines()
    assert lines[0].startswith("bad_create handle=0 class=org/apache/hadoop/hive/ql/exec/UDFArgumentException")
    assert "-frobnicate" in lines[0]
    assert lines[-1] == "done thrown=none"
    got = {int(l.split()[1]): np.float32(l.split()[2]) for l in lines if l.startswith("w ")}
    # the same rows (the driver's linear congruential stream) through the Python binding
    n, per = 1000, 6
    s = np.uint64(12345)
    f = np.zeros(n * per, np.int32); v = np.zeros(n * per, np.float32); y = np.zeros(n, np.float32)
    for i in range(n):
        acc = 0
        for k in range(per):
            s = (s * np.uint64(1664525) + np.uint64(1013904223)) & np.uint64(0xFFFFFFFF)
            f[i * per + k] = (int(s) >> 8) & 1023
            v[i * per + k] = np.float32(((int(s) >> 20) & 7) * 0.25 + 0.25)
            acc += int(f[i * per + k]) & 1
        y[i] = 1.0 if acc >= 3 else 0.0
    rp = np.arange(n + 1, dtype=np.int64) * per
    with hm.train_classifier(options) as u:
        u.process_batch(rp, f, v, y)
        feat, wt = u.forward()[:2]
    assert len(got) == len(feat) > 100
    assert all(got[int(j)] == w for j, w in zip(feat, wt))
