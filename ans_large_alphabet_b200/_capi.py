"""ctypes binding of include/b200ans.h (every exported symbol is declared here)."""
import ctypes as C
import os
import subprocess

import numpy as np

METHOD_INT, METHOD_FOLD, METHOD_MSB, METHOD_RFOLD = 0, 1, 2, 3
PATH_AUTO, PATH_GROUP, PATH_LPB = 0, 1, 2
HEADER_BYTES = 64

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
szp = C.POINTER(C.c_size_t)


class B200AnsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200ans error {code}: {msg}")
        self.code = code


class Stats(C.Structure):
    _fields_ = [
        ("ms_histogram", C.c_float), ("ms_model", C.c_float), ("ms_encode", C.c_float),
        ("ms_compact", C.c_float), ("ms_decode", C.c_float), ("ms_h2d", C.c_float), ("ms_d2h", C.c_float),
        ("launches", C.c_uint64), ("frame_size", C.c_uint64), ("n_symbols_model", C.c_uint32),
        ("n_blocks", C.c_uint32), ("prelude_bytes", C.c_uint64), ("payload_bytes", C.c_uint64),
        ("total_bytes", C.c_uint64), ("encode_kernel", C.c_char * 32), ("decode_kernel", C.c_char * 32),
        ("ms_allreduce", C.c_float), ("model_flags", C.c_uint32), ("model_cache_hits", C.c_uint64),
        ("peer_exchanges", C.c_uint64),
    ]


class Container(C.Structure):
    _fields_ = [
        ("method", C.c_int), ("fidelity", C.c_uint32), ("n_symbols", C.c_uint64), ("block_syms", C.c_uint32),
        ("n_blocks", C.c_uint32), ("prelude_bytes", C.c_uint64), ("payload_bytes", C.c_uint64),
        ("total_bytes", C.c_uint64),
    ]


# name -> (restype, argtypes); this is the complete export list of include/b200ans.h
SIGNATURES = {
    "b200ans_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_void_p]),
    "b200ans_destroy": (None, [C.c_void_p]),
    "b200ans_set_block_syms": (C.c_int, [C.c_void_p, C.c_uint32]),
    "b200ans_set_kernel_path": (C.c_int, [C.c_void_p, C.c_int]),
    "b200ans_suggest_block_syms": (C.c_uint32, [C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint32]),
    "b200ans_set_h_approx": (C.c_int, [C.c_void_p, C.c_uint32]),
    "b200ans_pa_compress_bound": (C.c_size_t, [C.c_void_p, C.c_int, C.c_uint32, C.c_size_t]),
    "b200ans_pa_compress": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_pa_decompress": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_pa_compress_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_pa_decompress_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_container_info": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(Container)]),
    "b200ans_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "b200ans_last_error": (C.c_char_p, [C.c_void_p]),
    "b200ans_strerror": (C.c_char_p, [C.c_int]),
    "b200ans_method_name": (C.c_char_p, [C.c_int, C.c_uint32, C.c_char_p, C.c_size_t]),
    "b200ans_compress_bound": (C.c_size_t, [C.c_void_p, C.c_int, C.c_size_t]),
    "b200ans_compress": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_decompress": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_compress_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_decompress_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_max_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_histogram_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_histogram_max_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_rfold_prepare": (C.c_int, [C.c_void_p, C.c_uint32, u64p, C.c_size_t]),
    "b200ans_model_create": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, u32p, C.c_size_t]),
    "b200ans_model_create_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t]),
    "b200ans_model_most_frequent": (C.c_int, [C.c_void_p, u32p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_uint64)]),
    "b200ans_pa_remap_compress": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, u8p, C.c_size_t, u32p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "b200ans_pa_remap_decompress": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, u32p, C.c_size_t, u8p, C.c_size_t]),
    "b200ans_pa_remap_compress_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "b200ans_pa_remap_decompress_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_byte_compress": (C.c_int, [C.c_void_p, u8p, C.c_size_t, u8p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "b200ans_byte_decompress": (C.c_int, [C.c_void_p, u8p, C.c_size_t, u8p, C.c_size_t]),
    "b200ans_byte_compress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "b200ans_byte_decompress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "b200ans_byte_adjust_freqs": (C.c_int, [u64p, C.POINTER(C.c_uint16)]),
    "b200ans_rfold_prepare_device": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_model_build": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, u64p, C.c_size_t]),
    "b200ans_model_prelude": (C.c_int, [C.c_void_p, u8p, C.c_size_t, szp]),
    "b200ans_model_nfreqs": (C.c_int, [C.c_void_p, u32p, C.c_size_t, szp, u64p]),
    "b200ans_model_load": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, u8p, C.c_size_t, szp]),
    "b200ans_encode_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_encode_blocks_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_block_stride": (C.c_int, [C.c_void_p, C.c_int, szp]),
    "b200ans_decode_blocks_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_decode_blocks_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_comm_unique_id": (C.c_int, [u8p]),
    "b200ans_comm_init": (C.c_int, [C.c_void_p, u8p, C.c_int, C.c_int]),
    "b200ans_comm_attach": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "b200ans_comm_ipc_export": (C.c_int, [C.c_void_p, u8p]),
    "b200ans_comm_ipc_attach": (C.c_int, [C.c_void_p, C.c_int, C.c_int, u8p]),
    "b200ans_comm_ipc_release": (C.c_int, [C.c_void_p]),
    "b200ans_comm_release": (C.c_int, [C.c_void_p]),
    "b200ans_allreduce_hist": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200ans_compress_sharded_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_compress_sharded": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, szp]),
    "b200ans_scale_freqs": (C.c_int, [u32p, u64p, C.c_size_t, u32p, C.c_size_t, C.c_int64, C.c_uint64, C.POINTER(C.c_int)]),
    "b200ans_adjust_freqs": (C.c_int, [u64p, C.c_size_t, C.c_uint32, C.c_int, C.c_uint32, u32p]),
    "b200ans_serialize_interp": (C.c_int, [u32p, C.c_size_t, C.c_uint64, u8p, C.c_size_t, szp]),
    "b200ans_load_interp": (C.c_int, [u8p, u32p, C.c_size_t, szp, szp]),
    "b200ans_interp_encode": (C.c_int, [u32p, u32p, C.c_size_t, C.c_size_t, szp]),
    "b200ans_interp_decode": (C.c_int, [u32p, u32p, C.c_size_t, C.c_size_t, szp]),
    "b200ans_fold_map": (C.c_uint32, [C.c_uint32, C.c_uint32, u32p]),
    "b200ans_fold_unmap": (C.c_uint32, [C.c_uint32, C.c_uint32, u32p]),
    "b200ans_msb_map": (C.c_uint32, [C.c_uint32, u32p]),
    "b200ans_msb_unmap": (C.c_uint32, [C.c_uint32, u32p]),
}


def library_path():
    return os.path.join(_HERE, "libb200ans.so")


def build_library(force=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> libb200ans.so (in-tree)."""
    if force and os.path.exists(library_path()):
        os.remove(library_path())
    subprocess.check_call(["make", "-C", _HERE, "libb200ans.so"])
    return library_path()


def load_library():
    """Load the CUDA library.  There is deliberately no fallback: a missing library is an error."""
    global _LIB
    if _LIB is None:
        path = library_path()
        if not os.path.exists(path):
            raise ImportError(
                f"{path} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()); "
                "b200ans has no CPU fallback")
        lib = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB


def comm_unique_id():
    """a fresh NCCL unique id (128 bytes, numpy uint8): rank 0 calls this and hands the bytes to every rank"""
    uid = np.zeros(128, dtype=np.uint8)
    rc = load_library().b200ans_comm_unique_id(uid.ctypes.data_as(u8p))
    if rc:
        raise B200AnsError(rc, "NCCL unavailable (libnccl.so.2)")
    return uid


def suggest_block_syms(n, lo=8192, hi=16384):
    """b200ans_suggest_block_syms without a context (148 SMs assumed): host arithmetic only"""
    return int(load_library().b200ans_suggest_block_syms(None, n, lo, hi))


def method_name(method, fidelity=0):
    buf = C.create_string_buffer(32)
    return load_library().b200ans_method_name(method, fidelity, buf, 32).decode()


def _p(a, t):
    return a.ctypes.data_as(t)


class _HostApi:
    """ans_util.hpp-shaped host functions (work without a GPU)."""

    def __init__(self):
        self.lib = load_library()

    def adjust_freqs(self, freqs, largest_sym, require_u16, h_approx=1):
        freqs = np.ascontiguousarray(freqs, dtype=np.uint64)
        out = np.zeros(largest_sym + 1, dtype=np.uint32)
        rc = self.lib.b200ans_adjust_freqs(_p(freqs, u64p), freqs.size, largest_sym, int(require_u16), h_approx, _p(out, u32p))
        if rc:
            raise B200AnsError(rc, "adjust_freqs")
        return out

    def byte_adjust_freqs(self, freqs):
        freqs = np.ascontiguousarray(freqs, dtype=np.uint64)
        assert freqs.size == 256
        out = np.zeros(256, dtype=np.uint16)
        rc = self.lib.b200ans_byte_adjust_freqs(_p(freqs, u64p), out.ctypes.data_as(C.POINTER(C.c_uint16)))
        if rc:
            raise B200AnsError(rc, "byte_adjust_freqs")
        return out

    def scale_freqs(self, freqs, mapping, frame, freq_sum=None):
        """scale_freqs (ans_util.hpp:77-95): (S, fits) for the counts `freqs`, symbols visited in `mapping` order"""
        freqs = np.ascontiguousarray(freqs, dtype=np.uint64)
        mapping = np.ascontiguousarray(mapping, dtype=np.uint32)
        out = np.zeros(freqs.size, dtype=np.uint32)
        fits = C.c_int()
        total = int(freqs.sum()) if freq_sum is None else int(freq_sum)
        rc = self.lib.b200ans_scale_freqs(_p(out, u32p), _p(freqs, u64p), freqs.size, _p(mapping, u32p), mapping.size, int(frame),
                                          total, C.byref(fits))
        if rc:
            raise B200AnsError(rc, "scale_freqs")
        return out, bool(fits.value)

    def serialize_interp(self, nfreqs, frame_size):
        nfreqs = np.ascontiguousarray(nfreqs, dtype=np.uint32)
        out = np.zeros(nfreqs.size * 8 + 64, dtype=np.uint8)
        nb = C.c_size_t()
        rc = self.lib.b200ans_serialize_interp(_p(nfreqs, u32p), nfreqs.size, frame_size, _p(out, u8p), out.size, C.byref(nb))
        if rc:
            raise B200AnsError(rc, "serialize_interp")
        return out[: nb.value].copy()

    def load_interp(self, blob):
        buf = np.zeros(len(blob) + 16, dtype=np.uint8)
        buf[: len(blob)] = blob
        cnt, used = C.c_size_t(), C.c_size_t()
        rc = self.lib.b200ans_load_interp(_p(buf, u8p), None, 0, C.byref(cnt), None)
        if rc:
            raise B200AnsError(rc, "load_interp")
        out = np.zeros(cnt.value, dtype=np.uint32)
        rc = self.lib.b200ans_load_interp(_p(buf, u8p), _p(out, u32p), out.size, C.byref(cnt), C.byref(used))
        if rc:
            raise B200AnsError(rc, "load_interp")
        return out, used.value

    def interp_encode(self, values, universe):
        values = np.ascontiguousarray(values, dtype=np.uint32)
        out = np.zeros(values.size * 2 + 16, dtype=np.uint32)
        nb = C.c_size_t()
        self.lib.b200ans_interp_encode(_p(out, u32p), _p(values, u32p), values.size, universe, C.byref(nb))
        return out[: nb.value // 4].copy()

    def interp_decode(self, words, n, universe):
        buf = np.zeros(len(words) + 4, dtype=np.uint32)
        buf[: len(words)] = words
        out = np.zeros(n, dtype=np.uint32)
        self.lib.b200ans_interp_decode(_p(buf, u32p), _p(out, u32p), n, universe, None)
        return out

    def container_info(self, blob):
        """header of a b200ans container as a dict (host only)"""
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        info = Container()
        rc = self.lib.b200ans_container_info(_p(blob, C.c_void_p), blob.size, C.byref(info))
        if rc:
            raise B200AnsError(rc, "not a b200ans container")
        return {k: getattr(info, k) for k, _ in Container._fields_}

    def fold_map(self, f, x):
        k = C.c_uint32()
        s = self.lib.b200ans_fold_map(f, int(x), C.byref(k))
        return s, k.value

    def fold_unmap(self, f, s):
        k = C.c_uint32()
        v = self.lib.b200ans_fold_unmap(f, int(s), C.byref(k))
        return v, k.value

    def msb_map(self, x):
        k = C.c_uint32()
        s = self.lib.b200ans_msb_map(int(x), C.byref(k))
        return s, k.value

    def msb_unmap(self, s):
        k = C.c_uint32()
        v = self.lib.b200ans_msb_unmap(int(s), C.byref(k))
        return v, k.value


def host_api():
    return _HostApi()


class Codec:
    """One b200ans context (one GPU, one stream).  Pointers may be numpy arrays (host entry
    points) or integers / torch tensors' data_ptr() (device entry points)."""

    def __init__(self, device=0, block_syms=8192, stream=None):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.b200ans_create(C.byref(h), device, stream)
        if rc:
            raise B200AnsError(rc, self.lib.b200ans_strerror(rc).decode())
        self.h = h
        self.set_block_syms(block_syms)

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200ans_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc:
            raise B200AnsError(rc, self.lib.b200ans_last_error(self.h).decode() or self.lib.b200ans_strerror(rc).decode())

    def set_block_syms(self, bs):
        self._check(self.lib.b200ans_set_block_syms(self.h, bs))
        self.block_syms = bs

    def suggest_block_syms(self, n, lo=8192, hi=16384):
        """block size in [lo, hi] whose blocks fill whole rounds of the lane-per-block coders on this GPU"""
        return int(self.lib.b200ans_suggest_block_syms(self.h, n, lo, hi))

    def set_h_approx(self, h):
        """adjust_freqs' H_approx: 1 = ANS / ANSmsb / ANSfold, other values = ANSsint<H> / ANSsmsb<H>"""
        self._check(self.lib.b200ans_set_h_approx(self.h, h))

    def set_kernel_path(self, path):
        """0 = auto, 1 = group kernels (4 lanes per block), 2 = lane-per-block kernels"""
        self._check(self.lib.b200ans_set_kernel_path(self.h, path))

    def stats(self):
        s = Stats()
        self._check(self.lib.b200ans_get_stats(self.h, C.byref(s)))
        return s

    def bound(self, method, n):
        return self.lib.b200ans_compress_bound(self.h, method, n)

    # ---- host buffers -------------------------------------------------------------------
    def compress(self, method, f, data, out=None):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        if out is None:
            out = np.empty(self.bound(method, data.size), dtype=np.uint8)
        nb = C.c_size_t()
        self._check(self.lib.b200ans_compress(self.h, method, f, out.ctypes.data, out.size, data.ctypes.data, data.size, C.byref(nb)))
        return out[: nb.value]

    def decompress(self, method, f, blob, n, out=None):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        if out is None:
            out = np.empty(n, dtype=np.uint32)
        self._check(self.lib.b200ans_decompress(self.h, method, f, out.ctypes.data, n, blob.ctypes.data, blob.size))
        return out

    # ---- pseudo-adaptive mode: one model per `model_syms` symbols (pseudo_adaptive.cpp) ----
    def pa_bound(self, method, model_syms, n):
        return self.lib.b200ans_pa_compress_bound(self.h, method, model_syms, n)

    def pa_compress(self, method, f, model_syms, data):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        out = np.empty(self.pa_bound(method, model_syms, data.size), dtype=np.uint8)
        nb = C.c_size_t()
        self._check(self.lib.b200ans_pa_compress(self.h, method, f, model_syms, out.ctypes.data, out.size, data.ctypes.data,
                                                 data.size, C.byref(nb)))
        return out[: nb.value]

    def pa_decompress(self, method, f, blob, n):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        out = np.empty(n, dtype=np.uint32)
        self._check(self.lib.b200ans_pa_decompress(self.h, method, f, out.ctypes.data, n, blob.ctypes.data, blob.size))
        return out

    def pa_compress_ptr(self, method, f, model_syms, dst_ptr, cap, src_ptr, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_pa_compress(self.h, method, f, model_syms, dst_ptr, cap, src_ptr, n, C.byref(nb)))
        return nb.value

    def pa_decompress_ptr(self, method, f, dst_ptr, n, src_ptr, nbytes):
        self._check(self.lib.b200ans_pa_decompress(self.h, method, f, dst_ptr, n, src_ptr, nbytes))

    def pa_compress_device(self, method, f, model_syms, d_dst, cap, d_src, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_pa_compress_device(self.h, method, f, model_syms, d_dst, cap, d_src, n, C.byref(nb)))
        return nb.value

    def pa_decompress_device(self, method, f, d_dst, n, d_src, nbytes):
        self._check(self.lib.b200ans_pa_decompress_device(self.h, method, f, d_dst, n, d_src, nbytes))

    def compress_ptr(self, method, f, dst_ptr, cap, src_ptr, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_compress(self.h, method, f, dst_ptr, cap, src_ptr, n, C.byref(nb)))
        return nb.value

    def decompress_ptr(self, method, f, dst_ptr, n, src_ptr, nbytes):
        self._check(self.lib.b200ans_decompress(self.h, method, f, dst_ptr, n, src_ptr, nbytes))

    # ---- device buffers (raw pointers) ------------------------------------------------------
    def compress_device(self, method, f, d_dst, cap, d_src, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_compress_device(self.h, method, f, d_dst, cap, d_src, n, C.byref(nb)))
        return nb.value

    def decompress_device(self, method, f, d_dst, n, d_src, nbytes):
        self._check(self.lib.b200ans_decompress_device(self.h, method, f, d_dst, n, d_src, nbytes))

    def decode_blocks_noalloc(self, d_payload, payload_bytes, d_block_bytes, n, d_dst):
        """container-layout blocks + the loaded model -> symbols, with no allocation (offsets in a static device array)"""
        self._check(self.lib.b200ans_decode_blocks_noalloc(self.h, d_payload, payload_bytes, d_block_bytes, n, d_dst))

    # ---- multi-GPU (one context per GPU; NCCL is opened by the library) ---------------------------
    def comm_init(self, unique_id, rank, world):
        """collective: ncclCommInitRank with the 128-byte id rank 0 got from comm_unique_id()"""
        uid = np.ascontiguousarray(unique_id, dtype=np.uint8)
        assert uid.size == 128
        self._check(self.lib.b200ans_comm_init(self.h, _p(uid, u8p), rank, world))

    def comm_attach(self, nccl_comm, rank, world):
        self._check(self.lib.b200ans_comm_attach(self.h, nccl_comm, rank, world))

    def comm_ipc_export(self):
        """this rank's exchange buffer for the histogram exchange over NVLink peer memory, as 64 opaque bytes"""
        h = np.zeros(64, dtype=np.uint8)
        self._check(self.lib.b200ans_comm_ipc_export(self.h, _p(h, u8p)))
        return h

    def comm_ipc_attach(self, rank, world, handles):
        """map the other ranks' exchange buffers (`handles`: world x 64 bytes in rank order)"""
        hs = np.ascontiguousarray(handles, dtype=np.uint8).reshape(-1)
        assert hs.size == 64 * world
        self._check(self.lib.b200ans_comm_ipc_attach(self.h, rank, world, _p(hs, u8p)))

    def comm_ipc_release(self):
        self._check(self.lib.b200ans_comm_ipc_release(self.h))

    def comm_release(self):
        self._check(self.lib.b200ans_comm_release(self.h))

    def allreduce_hist(self, d_bins, nbins, d_max=None):
        self._check(self.lib.b200ans_allreduce_hist(self.h, d_bins, nbins, d_max))

    def compress_sharded_device(self, method, f, d_dst, cap, d_src, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_compress_sharded_device(self.h, method, f, d_dst, cap, d_src, n, C.byref(nb)))
        return nb.value

    def compress_sharded_ptr(self, method, f, dst_ptr, cap, src_ptr, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_compress_sharded(self.h, method, f, dst_ptr, cap, src_ptr, n, C.byref(nb)))
        return nb.value

    def max_device(self, d_src, n, d_max):
        self._check(self.lib.b200ans_max_device(self.h, d_src, n, d_max))

    def histogram_device(self, method, f, d_src, n, d_bins, nbins):
        self._check(self.lib.b200ans_histogram_device(self.h, method, f, d_src, n, d_bins, nbins))

    def histogram_max_device(self, method, f, d_src, n, d_bins, nbins, d_max):
        self._check(self.lib.b200ans_histogram_max_device(self.h, method, f, d_src, n, d_bins, nbins, d_max))

    def rfold_prepare(self, f, raw_bins):
        raw_bins = np.ascontiguousarray(raw_bins, dtype=np.uint64)
        self._check(self.lib.b200ans_rfold_prepare(self.h, f, _p(raw_bins, u64p), raw_bins.size))

    def rfold_prepare_device(self, f, d_raw_bins, nvalues, d_folded_bins):
        self._check(self.lib.b200ans_rfold_prepare_device(self.h, f, d_raw_bins, nvalues, d_folded_bins))

    def pa_remap_compress(self, method, f, model_syms, data):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        out = np.zeros(self.pa_bound(method, model_syms, data.size) + data.size * 4 + (data.size // model_syms + 2) * 48 + 256, dtype=np.uint8)
        nb = C.c_size_t()
        self._check(self.lib.b200ans_pa_remap_compress(self.h, method, f, model_syms, _p(out, u8p), out.size, _p(data, u32p), data.size, C.byref(nb)))
        return out[: nb.value]

    def pa_remap_decompress(self, method, f, blob, n):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        out = np.zeros(n, dtype=np.uint32)
        self._check(self.lib.b200ans_pa_remap_decompress(self.h, method, f, _p(out, u32p), n, _p(blob, u8p), blob.size))
        return out

    def pa_remap_compress_device(self, method, f, model_syms, d_dst, cap, d_src, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_pa_remap_compress_device(self.h, method, f, model_syms, d_dst, cap, d_src, n, C.byref(nb)))
        return nb.value

    def pa_remap_decompress_device(self, method, f, d_dst, n, d_src, nbytes):
        self._check(self.lib.b200ans_pa_remap_decompress_device(self.h, method, f, d_dst, n, d_src, nbytes))

    def byte_compress(self, data):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        out = np.zeros(self.bound(4, data.size), dtype=np.uint8)
        nb = C.c_size_t()
        self._check(self.lib.b200ans_byte_compress(self.h, _p(out, u8p), out.size, _p(data, u8p), data.size, C.byref(nb)))
        return out[: nb.value]

    def byte_decompress(self, blob, n):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        out = np.zeros(n, dtype=np.uint8)
        self._check(self.lib.b200ans_byte_decompress(self.h, _p(out, u8p), n, _p(blob, u8p), blob.size))
        return out

    def model_create(self, method, f, data):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        self._check(self.lib.b200ans_model_create(self.h, method, f, _p(data, u32p), data.size))

    def model_create_device(self, method, f, d_src, n):
        self._check(self.lib.b200ans_model_create_device(self.h, method, f, d_src, n))

    def model_most_frequent(self):
        cnt, sigma = C.c_size_t(), C.c_uint64()
        self._check(self.lib.b200ans_model_most_frequent(self.h, None, 0, C.byref(cnt), C.byref(sigma)))
        out = np.zeros(cnt.value, dtype=np.uint32)
        if cnt.value:
            self._check(self.lib.b200ans_model_most_frequent(self.h, _p(out, u32p), out.size, C.byref(cnt), C.byref(sigma)))
        return out, sigma.value

    def model_build(self, method, f, bins):
        bins = np.ascontiguousarray(bins, dtype=np.uint64)
        self._check(self.lib.b200ans_model_build(self.h, method, f, _p(bins, u64p), bins.size))

    def model_prelude(self):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_model_prelude(self.h, None, 0, C.byref(nb)))
        out = np.zeros(nb.value, dtype=np.uint8)
        self._check(self.lib.b200ans_model_prelude(self.h, _p(out, u8p), out.size, C.byref(nb)))
        return out

    def model_nfreqs(self):
        cnt, fs = C.c_size_t(), C.c_uint64()
        self._check(self.lib.b200ans_model_nfreqs(self.h, None, 0, C.byref(cnt), C.byref(fs)))
        out = np.zeros(cnt.value, dtype=np.uint32)
        self._check(self.lib.b200ans_model_nfreqs(self.h, _p(out, u32p), out.size, C.byref(cnt), C.byref(fs)))
        return fs.value, out

    def model_load(self, method, f, prelude):
        prelude = np.ascontiguousarray(prelude, dtype=np.uint8)
        used = C.c_size_t()
        self._check(self.lib.b200ans_model_load(self.h, method, f, _p(prelude, u8p), prelude.size, C.byref(used)))
        return used.value

    def encode_device(self, d_dst, cap, d_src, n):
        nb = C.c_size_t()
        self._check(self.lib.b200ans_encode_device(self.h, d_dst, cap, d_src, n, C.byref(nb)))
        return nb.value

    def block_stride(self, method):
        s = C.c_size_t()
        self._check(self.lib.b200ans_block_stride(self.h, method, C.byref(s)))
        return s.value

    def encode_blocks_device(self, d_src, n, d_scratch, stride, d_block_bytes):
        self._check(self.lib.b200ans_encode_blocks_device(self.h, d_src, n, d_scratch, stride, d_block_bytes))

    def decode_blocks_device(self, d_payload, d_offsets, d_block_bytes, n, d_dst):
        self._check(self.lib.b200ans_decode_blocks_device(self.h, d_payload, d_offsets, d_block_bytes, n, d_dst))
