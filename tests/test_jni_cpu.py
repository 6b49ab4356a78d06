"""The JNI shim (bindings/jni/qegpu_jni.c) built against the JNI-layout stand-in header and a mock JNIEnv: no JVM is
needed to execute it.  Without a GPU every native that reaches the device fails -- which is the error path: status
!= 0 -> one pending IllegalStateException carrying qe_last_error(), the native still returns normally."""
import os
import re

import pytest

from jni_driver import PREFIX, ROOT, JavaException, QeGpu
from queryengine_b200.build import build


@pytest.fixture(scope="module")
def jni():
    build()
    return QeGpu()


def test_every_external_fun_has_its_native(jni):
    kt = open(os.path.join(ROOT, "bindings", "kotlin", "net", "jhorstmann", "queryengine", "gpu", "QeGpu.kt")).read()
    declared = set(re.findall(r"external fun (\w+)\(", kt))
    assert declared == set(jni.natives) and len(declared) == 31
    for name in declared:
        assert hasattr(jni.L, PREFIX + name)
    assert jni.L.mock_function_table_slots() == 234           # JNI specification: reserved0 .. GetModule


def test_failure_becomes_one_pending_illegal_state_exception(jni):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the error path is covered by tests/test_gpu_jni.py")
    with pytest.raises(JavaException) as ei:
        jni.call("ctxCreate", 0)
    assert ei.value.cls == "java/lang/IllegalStateException" and "no CPU fallback" in ei.value.message
    # arrays are released on the error path too
    words = jni.int_array([1, 2, 3])
    with pytest.raises(JavaException) as ei:
        jni.call("programCreate", 0, words)
    assert "ctx is NULL" in ei.value.message and jni.outstanding == 0
    jni.free(words)
