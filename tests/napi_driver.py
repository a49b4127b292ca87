"""ctypes driver for napi/apm_b200_addon.c linked against the fake Node-API of tests/fake_napi.c."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
U8, U32, I32, F64, BIGI64, BIGU64 = 1, 6, 5, 8, 9, 10  # napi_typedarray_type


def load():
    so = os.path.join(ROOT, "tests", "_fake_napi_addon.so")
    srcs = [os.path.join(ROOT, "napi", "apm_b200_addon.c"), os.path.join(ROOT, "tests", "fake_napi.c")]
    if not os.path.exists(so) or any(os.path.getmtime(so) < os.path.getmtime(s) for s in srcs):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "napi"), "../tests/_fake_napi_addon.so"],
                              stdout=subprocess.DEVNULL)
    L = C.CDLL(so)
    for name in ("fk_env", "fk_object", "fk_init_module", "fk_get", "fk_new", "fk_call_method", "fk_call_function", "fk_buffer_data"):
        getattr(L, name).restype = C.c_void_p
    L.fk_number.restype = C.c_void_p
    L.fk_number.argtypes = [C.c_double]
    L.fk_string.restype = C.c_void_p
    L.fk_string.argtypes = [C.c_char_p]
    L.fk_array.restype = C.c_void_p
    L.fk_array.argtypes = [C.c_void_p, C.c_uint32]
    L.fk_array_of.restype = C.c_void_p
    L.fk_array_of.argtypes = [C.c_void_p, C.c_uint32]
    L.fk_typedarray.restype = C.c_void_p
    L.fk_typedarray.argtypes = [C.c_int, C.c_void_p, C.c_size_t]
    L.fk_set.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
    L.fk_get.argtypes = [C.c_void_p, C.c_char_p]
    L.fk_new.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    L.fk_call_method.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_void_p]
    L.fk_call_function.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    L.fk_kind_of.argtypes = [C.c_void_p]
    L.fk_number_value.argtypes = [C.c_void_p]
    L.fk_number_value.restype = C.c_double
    L.fk_buffer_data.argtypes = [C.c_void_p]
    L.fk_buffer_len.argtypes = [C.c_void_p]
    L.fk_buffer_len.restype = C.c_size_t
    L.fk_exception.restype = C.c_char_p
    return L


class JsError(Exception):
    pass


class FakeNode:
    """Just enough of a JS runtime to `require` the addon and call its methods."""

    def __init__(self):
        self.L = load()
        self.exports = self.L.fk_init_module()
        self._keep = []

    def obj(self, d: dict):
        o = self.L.fk_object()
        for k, v in d.items():
            if isinstance(v, (list, tuple)):
                a = np.ascontiguousarray(v, dtype=np.float64)
                self._keep.append(a)
                self.L.fk_set(o, k.encode(), self.L.fk_array(a.ctypes.data, a.size))
            else:
                self.L.fk_set(o, k.encode(), self.L.fk_number(float(v)))
        return o

    def typed(self, arr: np.ndarray):
        kind = {np.dtype(np.uint8): U8, np.dtype(np.uint32): U32, np.dtype(np.int32): I32, np.dtype(np.int64): BIGI64,
                np.dtype(np.uint64): BIGU64, np.dtype(np.float64): F64}[arr.dtype]
        a = np.ascontiguousarray(arr)
        self._keep.append(a)
        return self.L.fk_typedarray(kind, a.ctypes.data, a.size)

    def _args(self, args):
        arr = (C.c_void_p * max(len(args), 1))(*args)
        self._keep.append(arr)
        return arr

    def new(self, cls_name: str, *args):
        cls = self.L.fk_get(self.exports, cls_name.encode())
        inst = self.L.fk_new(cls, len(args), self._args(args))
        if not inst:
            raise JsError(self.L.fk_exception().decode())
        return inst

    def call(self, inst, method: str, *args):
        r = self.L.fk_call_method(inst, method.encode(), len(args), self._args(args))
        if not r:
            raise JsError(self.L.fk_exception().decode())
        return r

    def fn(self, name: str, *args):
        """a module-level function of the addon (exports.name(...))"""
        r = self.L.fk_call_function(self.L.fk_get(self.exports, name.encode()), len(args), self._args(args))
        if not r:
            raise JsError(self.L.fk_exception().decode())
        return r

    def string_value(self, v) -> str:
        self.L.fk_string_value.restype = C.c_char_p
        self.L.fk_string_value.argtypes = [C.c_void_p]
        return self.L.fk_string_value(v).decode()

    def number(self, x):
        return self.L.fk_number(float(x))

    def string(self, s):
        return self.L.fk_string(s.encode())

    def strings(self, items):
        vals = self._args([self.string(x) for x in items])
        return self.L.fk_array_of(vals, len(items))

    def numbers(self, items):
        a = np.ascontiguousarray(items, dtype=np.float64)
        self._keep.append(a)
        return self.L.fk_array(a.ctypes.data, a.size)

    def prop(self, o, name):
        return self.L.fk_get(o, name.encode())

    def buffer(self, v, dtype) -> np.ndarray:
        n = self.L.fk_buffer_len(v)
        if n == 0:
            return np.zeros(0, dtype=dtype)
        raw = (C.c_char * n).from_address(self.L.fk_buffer_data(v))
        return np.frombuffer(raw, dtype=dtype).copy()
