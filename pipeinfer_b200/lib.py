"""ctypes mirror of include/pipeinfer_b200.h (operator level).  Fails loudly when the CUDA library is missing."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libpipeinfer_b200.so")

TYPE_Q4_0, TYPE_Q8_0, TYPE_Q4_K = 2, 8, 12
TYPE_Q3_K = 11        # matmul + embedding rows (Q3_K_M files); activations as Q4_K
TYPE_Q5_K = 13
TYPE_Q6_K = 14        # matmul only; multiplies with the same activations as Q4_K
TYPES = {"q4_0": TYPE_Q4_0, "q8_0": TYPE_Q8_0, "q4_k": TYPE_Q4_K}
BLOCK = {TYPE_Q4_0: (32, 18), TYPE_Q8_0: (32, 34), TYPE_Q4_K: (256, 144), TYPE_Q5_K: (256, 176), TYPE_Q6_K: (256, 210), TYPE_Q3_K: (256, 110)}
ACT_BLOCK = {TYPE_Q4_0: (32, 34), TYPE_Q8_0: (32, 34), TYPE_Q4_K: (256, 292)}


class LibraryMissing(RuntimeError):
    pass


class PibError(RuntimeError):
    pass


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Lib:
    def __init__(self, path=SO_PATH):
        if not os.path.exists(path):
            raise LibraryMissing("%s not built: run `python __graft_entry__.py build` (needs nvcc); "
                                 "there is no CPU fallback" % path)
        L = self.L = C.CDLL(path)
        L.pib_last_error.restype = C.c_char_p
        L.pib_init.argtypes = [C.c_int]
        L.pib_dev_alloc.restype = C.c_void_p
        L.pib_dev_alloc.argtypes = [C.c_size_t]
        L.pib_dev_free.argtypes = [C.c_void_p]
        L.pib_memcpy_h2d.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.pib_memcpy_d2h.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.pib_stream_sync.argtypes = [C.c_void_p]
        L.pib_weight_upload.restype = C.c_void_p
        L.pib_weight_upload.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64]
        L.pib_weight_from_device.restype = C.c_void_p
        L.pib_weight_from_device.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64]
        L.pib_weight_free.argtypes = [C.c_void_p]
        L.pib_weight_bytes.restype = C.c_size_t
        L.pib_weight_bytes.argtypes = [C.c_void_p]
        L.pib_weight_row_blocks.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        L.pib_act_bytes.restype = C.c_size_t
        L.pib_act_bytes.argtypes = [C.c_int, C.c_int64, C.c_int]
        L.pib_quantize_act.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.pib_norm_quantize_act.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int64,
                                            C.c_void_p, C.c_void_p, C.c_void_p]
        L.pib_act_to_blocks.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_void_p]
        L.pib_gemv.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.pib_gemv_q4k_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.pib_mul_mat_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.pib_get_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]

    def check(self, rc):
        if rc != 0:
            raise PibError("libpipeinfer_b200 error %d: %s" % (rc, self.L.pib_last_error().decode()))

    def init(self, device=0):
        self.check(self.L.pib_init(device))

    # ---- device buffers ----
    def to_device(self, arr):
        arr = np.ascontiguousarray(arr)
        d = self.L.pib_dev_alloc(arr.nbytes)
        if not d:
            raise PibError(self.L.pib_last_error().decode())
        self.check(self.L.pib_memcpy_h2d(d, _p(arr), arr.nbytes, None))
        self.check(self.L.pib_stream_sync(None))
        return d

    def alloc(self, nbytes):
        d = self.L.pib_dev_alloc(nbytes)
        if not d:
            raise PibError(self.L.pib_last_error().decode())
        return d

    def to_host(self, dev, shape, dtype):
        out = np.empty(shape, dtype=dtype)
        self.check(self.L.pib_memcpy_d2h(_p(out), dev, out.nbytes, None))
        self.check(self.L.pib_stream_sync(None))
        return out

    def free(self, dev):
        self.L.pib_dev_free(dev)

    # ---- operators on numpy arrays (host in, host out) ----
    def weight_upload(self, wtype, blocks, K):
        blocks = np.ascontiguousarray(blocks)
        N = blocks.shape[0]
        w = self.L.pib_weight_upload(wtype, _p(blocks), N, K)
        if not w:
            raise PibError(self.L.pib_last_error().decode())
        return w

    def weight_rows(self, w, wtype, N, K):
        be, bb = BLOCK[wtype]
        out = np.zeros((N, K // be * bb), dtype=np.uint8)
        for r in range(N):
            self.check(self.L.pib_weight_row_blocks(w, r, _p(out[r])))
        return out

    def quantize_act(self, wtype, x, norm_w=None, eps=0.0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        b, K = x.shape
        xd = self.to_device(x)
        act = self.alloc(self.L.pib_act_bytes(wtype, K, b))
        if norm_w is None:
            self.check(self.L.pib_quantize_act(wtype, xd, b, K, act, None))
        else:
            wd = self.to_device(np.ascontiguousarray(norm_w, dtype=np.float32))
            self.check(self.L.pib_norm_quantize_act(wtype, 1, xd, wd, eps, b, K, act, None, None))
            self.free(wd)
        be, bb = ACT_BLOCK[wtype]
        out = np.zeros((b, K // be * bb), dtype=np.uint8)
        self.check(self.L.pib_act_to_blocks(wtype, act, b, K, _p(out)))
        self.free(xd)
        self.free(act)
        return out

    def mul_mat(self, w, x, N, variant=0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        b = x.shape[0]
        y = np.empty((b, N), dtype=np.float32)
        self.check(self.L.pib_mul_mat_host(w, _p(x), b, _p(y), variant))
        return y

    def get_rows(self, w, ids, K):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        idd = self.to_device(ids)
        out = self.alloc(len(ids) * K * 4)
        self.check(self.L.pib_get_rows(w, idd, len(ids), out, None))
        res = self.to_host(out, (len(ids), K), np.float32)
        self.free(idd)
        self.free(out)
        return res


_lib = None


def load():
    global _lib
    if _lib is None:
        _lib = Lib()
    return _lib
