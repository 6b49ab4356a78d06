"""Drives the JNI natives of bindings/jni/qegpu_jni.c through a mock JNIEnv (tests/jni_mock.c): what the Kotlin object
QeGpu (bindings/kotlin/.../gpu/QeGpu.kt) would call on a JVM, called from ctypes instead.  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "_build")
LIB = os.path.join(BUILD, "libqegpu_jni_mock.so")
PREFIX = "Java_net_jhorstmann_queryengine_gpu_QeGpu_"


def build() -> str:
    """gcc: the shim + the mock env, linked against the in-tree libqegpu.so (what a maintainer does with a real JDK's
    jni.h instead of the stand-in header)."""
    os.makedirs(BUILD, exist_ok=True)
    srcs = [os.path.join(ROOT, "bindings", "jni", "qegpu_jni.c"), os.path.join(ROOT, "tests", "jni_mock.c")]
    newest = max(os.path.getmtime(p) for p in srcs + [os.path.join(ROOT, "include", "qegpu.h"),
                                                       os.path.join(ROOT, "bindings", "jni", "jni_stub", "jni.h")])
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
        libdir = os.path.join(ROOT, "queryengine_b200")
        subprocess.run(["gcc", "-shared", "-fPIC", "-O1", "-Wall", "-Werror", "-Wno-unused-parameter", "-fvisibility=hidden",
                        "-I", os.path.join(ROOT, "bindings", "jni", "jni_stub"), "-I", os.path.join(ROOT, "include")] + srcs +
                       ["-L", libdir, "-lqegpu", "-Wl,-rpath," + libdir, "-o", LIB], check=True)
    return LIB


class JavaException(Exception):
    """What the JVM would throw in the caller once the native method has returned."""

    def __init__(self, cls: str, message: str):
        super().__init__(f"{cls}: {message}")
        self.cls, self.message = cls, message


class QeGpu:
    """The `external fun`s of the Kotlin object QeGpu, same names and argument order."""

    def __init__(self):
        self.L = C.CDLL(build())
        L = self.L
        L.mock_env.restype = C.c_void_p
        for f in ("mock_new_int_array", "mock_new_long_array", "mock_new_boolean_array"):
            getattr(L, f).restype = C.c_void_p
            getattr(L, f).argtypes = [C.c_void_p, C.c_int32]
        L.mock_new_direct_buffer.restype = C.c_void_p
        L.mock_new_direct_buffer.argtypes = [C.c_void_p, C.c_int64]
        L.mock_free_object.argtypes = [C.c_void_p]
        L.mock_boolean_array_get.argtypes = [C.c_void_p, C.c_int32]
        L.mock_exception_class.restype = C.c_char_p
        L.mock_exception_message.restype = C.c_char_p
        self.env = C.c_void_p(L.mock_env())
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        sig = {
            "ctxCreate": (i64, [i32]), "ctxDestroy": (None, [i64]),
            "programCreate": (i64, [i64, vp]), "programDestroy": (None, [i64]),
            "tableCreate": (i64, [i64, vp, vp, i64]), "tableUpload": (None, [i64, i32, vp, vp, i64, i64]), "tableDestroy": (None, [i64]),
            "pipelineRun": (i64, [i64, i64, i64, vp, vp, i64]),
            "groupByRun": (i64, [i64, i64, i64, vp, vp, vp, i64, i64]),
            "resultNumRows": (i64, [i64]), "resultNumColumns": (i32, [i64]), "resultColumnType": (i32, [i64, i32]),
            "resultColumn": (None, [i64, i32, vp, vp]), "resultAgg": (C.c_double, [i64, i32, vp]), "resultDestroy": (None, [i64]),
            "pipelineRunFlags": (i64, [i64, i64, i64, vp, vp, i32, i64]), "resultSelection": (None, [i64, vp]),
            "mgpuCreate": (i64, [vp]), "mgpuDestroy": (None, [i64]), "mgpuCtx": (i64, [i64, i32]),
            "mtableCreate": (i64, [i64, vp, vp, i64]), "mtableUpload": (None, [i64, i32, vp, vp]), "mtableDestroy": (None, [i64]),
            "mgpuPipelineRun": (i64, [i64, i64, i64, vp, vp, i32]), "mgpuGroupByRun": (i64, [i64, i64, i64, vp, vp, vp, i64]),
            "mresultNumRows": (i64, [i64]), "mresultNumColumns": (i32, [i64]), "mresultColumn": (None, [i64, i32, vp, vp]),
            "mresultSelection": (None, [i64, vp]), "mresultAgg": (C.c_double, [i64, i32, vp]), "mresultDestroy": (None, [i64]),
        }
        self.natives = sorted(sig)
        for name, (res, args) in sig.items():
            f = getattr(L, PREFIX + name)
            f.restype = res
            f.argtypes = [vp, vp] + args                     # JNIEnv *, jclass, ...
            setattr(self, "_" + name, f)

    # ---- Java-side objects
    def int_array(self, values):
        a = np.ascontiguousarray(values, dtype=np.int32)
        return self.L.mock_new_int_array(a.ctypes.data, a.size)

    def long_array(self, values):
        a = np.ascontiguousarray(values, dtype=np.int64)
        return self.L.mock_new_long_array(a.ctypes.data, a.size)

    def boolean_array(self, values):
        a = np.ascontiguousarray(values, dtype=np.uint8)
        return self.L.mock_new_boolean_array(a.ctypes.data, a.size)

    def direct_buffer(self, array: np.ndarray):
        """ByteBuffer.allocateDirect(...) over the array's memory (the caller keeps the array alive)."""
        return self.L.mock_new_direct_buffer(array.ctypes.data, array.nbytes)

    def free(self, *objs):
        for o in objs:
            if o:
                self.L.mock_free_object(o)

    def call(self, name, *args):
        """Invoke a native; afterwards do what the JVM does: a pending exception is thrown in the caller."""
        r = getattr(self, "_" + name)(self.env, None, *args)
        if self.L.mock_exception_pending():
            cls, msg = self.L.mock_exception_class().decode(), self.L.mock_exception_message().decode(errors="replace")
            self.L.mock_exception_clear()
            raise JavaException(cls, msg)
        return r

    @property
    def outstanding(self) -> int:
        return self.L.mock_outstanding_array_elements()
