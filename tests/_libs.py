"""ctypes loaders shared by the tests (and by bench.py's cpu_baseline leg).

Three libraries:
  * ``product()``  -- ans_large_alphabet_b200/libb200ans.so, the C-ABI of include/b200ans.h
  * ``oracle()``   -- oracle/_ref/libansoracle.so, the plain-C restatement (test infrastructure)
  * ``reference()``-- oracle/_ref/libansref.so, the unmodified reference headers behind a shim
                      (test infrastructure; built only where /root/reference exists, but the
                      built file travels to the GPU box).
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle", "_ref")

M_INT, M_FOLD, M_MSB, M_RFOLD = 0, 1, 2, 3
METHOD_NAMES = {M_INT: "ANS", M_FOLD: "ANSfold", M_MSB: "ANSmsb", M_RFOLD: "ANSrfold"}

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)


def ptr(a, t):
    return a.ctypes.data_as(t)


class _CpuCodec:
    """Common python face of the oracle ("orc_") and the reference shim ("ref_")."""

    def __init__(self, path, prefix):
        self.lib = C.CDLL(path)
        self.p = prefix
        L = self.lib
        g = lambda n: getattr(L, prefix + n)
        g("compress").restype = C.c_long
        g("compress").argtypes = [C.c_int, C.c_uint32, u8p, C.c_size_t, u32p, C.c_size_t]
        g("decompress").restype = C.c_int
        g("decompress").argtypes = [C.c_int, C.c_uint32, u32p, C.c_size_t, u8p, C.c_size_t]
        g("adjust_freqs").restype = C.c_int
        g("adjust_freqs").argtypes = [u64p, C.c_size_t, C.c_uint32, C.c_int, C.c_uint32, u32p]
        g("serialize_interp").restype = C.c_long
        g("serialize_interp").argtypes = [u32p, C.c_size_t, C.c_uint64, u8p]
        g("load_interp").restype = C.c_long
        g("load_interp").argtypes = [u8p, u32p]
        g("interp_encode").restype = C.c_long
        g("interp_encode").argtypes = [u32p, u32p, C.c_size_t, C.c_size_t]
        g("interp_decode").restype = C.c_long
        g("interp_decode").argtypes = [u32p, u32p, C.c_size_t, C.c_size_t]
        g("model_create").restype = C.c_void_p
        g("model_create").argtypes = [C.c_int, C.c_uint32, u32p, C.c_size_t]
        g("model_load").restype = C.c_void_p
        g("model_load").argtypes = [C.c_int, C.c_uint32, u8p]
        g("model_info").restype = C.c_long
        g("model_info").argtypes = [C.c_void_p, u64p, u32p]
        g("model_serialize").restype = C.c_long
        g("model_serialize").argtypes = [C.c_void_p, u8p]
        g("model_rfold_info").restype = C.c_long
        g("model_rfold_info").argtypes = [C.c_void_p, u64p, u32p]
        g("encode_block").restype = C.c_long
        g("encode_block").argtypes = [C.c_void_p, u32p, C.c_size_t, u8p]
        g("decode_block").restype = C.c_int
        g("decode_block").argtypes = [C.c_void_p, u8p, C.c_size_t, u32p, C.c_size_t]
        g("model_free").restype = None
        g("model_free").argtypes = [C.c_void_p]
        g("time").restype = C.c_int
        g("time").argtypes = [C.c_int, C.c_uint32, u32p, C.c_size_t, C.c_int,
                              C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_size_t)]

        if prefix == "ref_" and hasattr(L, "ref_compress_h"):  # ANSsint<H> / ANSsmsb<H>: reference shim only
            L.ref_compress_h.restype = C.c_long
            L.ref_compress_h.argtypes = [C.c_int, C.c_uint32, u8p, C.c_size_t, u32p, C.c_size_t]
            L.ref_decompress_h.restype = C.c_int
            L.ref_decompress_h.argtypes = [C.c_int, u32p, C.c_size_t, u8p, C.c_size_t]

    def _f(self, name):
        return getattr(self.lib, self.p + name)

    def compress_h(self, method, h_approx, data):
        """ans_sint_compress<H> / ans_smsb_compress<H> of the unmodified reference"""
        data = np.ascontiguousarray(data, dtype=np.uint32)
        out = np.zeros(data.size * 8 + (1 << 16) + (8 << 20) * (method == M_INT), dtype=np.uint8)
        nb = self.lib.ref_compress_h(method, h_approx, ptr(out, u8p), out.size, ptr(data, u32p), data.size)
        assert nb >= 0
        return out[:nb].copy()

    def decompress_h(self, method, blob, n):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        buf = np.zeros(blob.size + 16, dtype=np.uint8)
        buf[: blob.size] = blob
        out = np.zeros(n + 8, dtype=np.uint32)
        assert self.lib.ref_decompress_h(method, ptr(out, u32p), n, ptr(buf, u8p), blob.size) == 0
        return out[:n].copy()

    # ---- whole stream -----------------------------------------------------------------
    def compress(self, method, f, data):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        out = np.zeros(data.size * 8 + (1 << 16) + (8 << 20) * (method in (M_INT, M_RFOLD)), dtype=np.uint8)
        nb = self._f("compress")(method, f, ptr(out, u8p), out.size, ptr(data, u32p), data.size)
        assert nb >= 0
        return out[:nb].copy()

    def decompress(self, method, f, blob, n):
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        # the reference decoder over-reads up to 3 bytes past the cursor: keep slack
        buf = np.zeros(blob.size + 16, dtype=np.uint8)
        buf[: blob.size] = blob
        out = np.zeros(n + 8, dtype=np.uint32)
        rc = self._f("decompress")(method, f, ptr(out, u32p), n, ptr(buf, u8p), blob.size)
        assert rc == 0
        return out[:n].copy()

    def pa_remap_block(self, method, f, data):
        """one block of src/pseudo_adaptive.cpp:83-125 -> (bytes, alphabet bytes, ranks); None when the driver's
        32-bit running sums of the block alphabet would wrap"""
        fn = self._f("pa_remap_block")
        fn.restype = C.c_long
        fn.argtypes = [C.c_int, C.c_uint32, u32p, C.c_size_t, u8p, C.c_size_t, u32p, C.POINTER(C.c_size_t)]
        data = np.ascontiguousarray(data, dtype=np.uint32)
        out = np.zeros(data.size * 16 + (1 << 16) + (8 << 20) * (method == M_INT), dtype=np.uint8)
        ranks = np.zeros(data.size, dtype=np.uint32)
        ab = C.c_size_t()
        nb = fn(method, f, ptr(data, u32p), data.size, ptr(out, u8p), out.size, ptr(ranks, u32p), C.byref(ab))
        if nb == -2:
            return None
        assert nb >= 0, nb
        return out[:nb].copy(), ab.value, ranks

    # ---- model utilities ----------------------------------------------------------------
    def adjust_freqs(self, freqs, largest_sym, require_u16, h_approx=1):
        freqs = np.ascontiguousarray(freqs, dtype=np.uint64)
        out = np.zeros(largest_sym + 1, dtype=np.uint32)
        self._f("adjust_freqs")(ptr(freqs, u64p), freqs.size, largest_sym, int(require_u16), h_approx,
                                ptr(out, u32p))
        return out

    def scale_freqs(self, freqs, mapping, frame, freq_sum=None):
        """reference shim only: scale_freqs (ans_util.hpp:77-95) -> (S, fits)"""
        fn = self.lib.ref_scale_freqs
        fn.restype = C.c_int
        fn.argtypes = [u32p, u64p, C.c_size_t, u32p, C.c_size_t, C.c_int64, C.c_uint64]
        freqs = np.ascontiguousarray(freqs, dtype=np.uint64)
        mapping = np.ascontiguousarray(mapping, dtype=np.uint32)
        out = np.zeros(freqs.size, dtype=np.uint32)
        failed = fn(ptr(out, u32p), ptr(freqs, u64p), freqs.size, ptr(mapping, u32p), mapping.size, int(frame),
                    int(freqs.sum()) if freq_sum is None else int(freq_sum))
        return out, not failed

    def serialize_interp(self, nfreqs, frame_size):
        nfreqs = np.ascontiguousarray(nfreqs, dtype=np.uint32)
        out = np.zeros(nfreqs.size * 8 + 64, dtype=np.uint8)
        nb = self._f("serialize_interp")(ptr(nfreqs, u32p), nfreqs.size, frame_size, ptr(out, u8p))
        return out[:nb].copy()

    def load_interp(self, blob):
        buf = np.zeros(len(blob) + 16, dtype=np.uint8)
        buf[: len(blob)] = blob
        cnt = self._f("load_interp")(ptr(buf, u8p), None) if self.p == "orc_" else None
        if cnt is None:
            # the reference shim has no count query that avoids a decode; decode twice
            cnt = self._f("load_interp")(ptr(buf, u8p), None)
        out = np.zeros(cnt, dtype=np.uint32)
        self._f("load_interp")(ptr(buf, u8p), ptr(out, u32p))
        return out

    def interp_encode(self, values, universe):
        values = np.ascontiguousarray(values, dtype=np.uint32)
        out = np.zeros(values.size * 2 + 16, dtype=np.uint32)
        nb = self._f("interp_encode")(ptr(out, u32p), ptr(values, u32p), values.size, universe)
        return out[: nb // 4].copy()

    def interp_decode(self, words, n, universe):
        buf = np.zeros(len(words) + 4, dtype=np.uint32)
        buf[: len(words)] = words
        out = np.zeros(n, dtype=np.uint32)
        self._f("interp_decode")(ptr(buf, u32p), ptr(out, u32p), n, universe)
        return out

    # ---- models / per-block -------------------------------------------------------------
    def model_create(self, method, f, data):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        h = self._f("model_create")(method, f, ptr(data, u32p), data.size)
        assert h
        return _Model(self, h)

    def model_from_hist(self, method, f, bins):
        """reference shim only: the encoder model create() would build for an input with these counts (2^30 / 2^32-symbol
        jobs: the counts exist, a CPU counting pass would take minutes); ANS / ANSmsb / ANSfold-f"""
        fn = self._f("model_from_hist")
        fn.restype = C.c_void_p
        fn.argtypes = [C.c_int, C.c_uint32, u64p, C.c_size_t]
        bins = np.ascontiguousarray(bins, dtype=np.uint64)
        h = fn(method, f, ptr(bins, u64p), bins.size)
        assert h
        return _Model(self, h)

    def model_load(self, method, f, prelude):
        buf = np.zeros(len(prelude) + 16, dtype=np.uint8)
        buf[: len(prelude)] = prelude
        h = self._f("model_load")(method, f, ptr(buf, u8p))
        assert h
        return _Model(self, h)

    def time(self, method, f, data, runs=1, block=0):
        """(encode ns, decode ns, bytes); block > 0: every `block` symbols compressed on their own (pseudo_adaptive.cpp)"""
        data = np.ascontiguousarray(data, dtype=np.uint32)
        e, d, b = C.c_double(), C.c_double(), C.c_size_t()
        if block:
            if hasattr(self.lib, self.p + "time_blocks"):
                fn = self._f("time_blocks")
                fn.restype = C.c_int
                fn.argtypes = [C.c_int, C.c_uint32, u32p, C.c_size_t, C.c_size_t, C.c_int,
                               C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_size_t)]
                rc = fn(method, f, ptr(data, u32p), data.size, block, runs, C.byref(e), C.byref(d), C.byref(b))
                assert rc == 0, rc
                return e.value, d.value, b.value
            te = td = 0.0
            tb = 0
            for i in range(0, data.size, block):
                ei, di, bi = self.time(method, f, data[i : i + block], runs)
                te, td, tb = te + ei, td + di, tb + bi
            return te, td, tb
        rc = self._f("time")(method, f, ptr(data, u32p), data.size, runs, C.byref(e), C.byref(d), C.byref(b))
        assert rc == 0, rc
        return e.value, d.value, b.value


class _Model:
    def __init__(self, codec, handle):
        self.c, self.h = codec, handle

    def info(self):
        fs = C.c_uint64()
        cnt = self.c._f("model_info")(self.h, C.byref(fs), None)
        nf = np.zeros(cnt, dtype=np.uint32)
        self.c._f("model_info")(self.h, C.byref(fs), ptr(nf, u32p))
        return fs.value, nf

    def rfold_info(self):
        sg = C.c_uint64()
        cnt = self.c._f("model_rfold_info")(self.h, C.byref(sg), None)
        mf = np.zeros(max(cnt, 0), dtype=np.uint32)
        if cnt > 0:
            self.c._f("model_rfold_info")(self.h, C.byref(sg), ptr(mf, u32p))
        return sg.value, mf

    def serialize(self):
        out = np.zeros(64 << 20, dtype=np.uint8) if False else np.zeros(self.info()[1].size * 8 + (1 << 16), dtype=np.uint8)
        nb = self.c._f("model_serialize")(self.h, ptr(out, u8p))
        return out[:nb].copy()

    def encode_block(self, data):
        data = np.ascontiguousarray(data, dtype=np.uint32)
        out = np.zeros(data.size * 8 + 64, dtype=np.uint8)
        nb = self.c._f("encode_block")(self.h, ptr(data, u32p), data.size, ptr(out, u8p))
        return out[:nb].copy()

    def decode_block(self, payload, n):
        buf = np.zeros(len(payload) + 16, dtype=np.uint8)
        buf[8 : 8 + len(payload)] = payload  # slack on both sides (decoder reads backwards)
        out = np.zeros(n + 8, dtype=np.uint32)
        view = buf[8:]
        self.c._f("decode_block")(self.h, ptr(view, u8p), len(payload), ptr(out, u32p), n)
        return out[:n].copy()

    def free(self):
        if self.h:
            self.c._f("model_free")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


_cache = {}


def oracle():
    if "o" not in _cache:
        _cache["o"] = _CpuCodec(os.path.join(ORACLE_DIR, "libansoracle.so"), "orc_")
        L = _cache["o"].lib
        L.orc_fold_mapping.restype = C.c_uint32
        L.orc_fold_mapping.argtypes = [C.c_uint32, C.c_uint32, u32p]
        L.orc_fold_undo.argtypes = [C.c_uint32, C.c_uint32, u32p, u32p]
        L.orc_msb_mapping.restype = C.c_uint32
        L.orc_msb_mapping.argtypes = [C.c_uint32, u32p]
        L.orc_msb_undo.argtypes = [C.c_uint32, u32p, u32p]
    return _cache["o"]


def have_reference():
    return os.path.exists(os.path.join(ORACLE_DIR, "libansref.so"))


def reference():
    if "r" not in _cache:
        _cache["r"] = _CpuCodec(os.path.join(ORACLE_DIR, "libansref.so"), "ref_")
        L = _cache["r"].lib
        L.ref_fold_mapping.argtypes = [C.c_uint32, C.c_uint32, u32p]
        L.ref_fold_undo.argtypes = [C.c_uint32, C.c_uint32, u32p, u32p]
        L.ref_msb_mapping.restype = C.c_uint32
        L.ref_msb_mapping.argtypes = [C.c_uint32]
        L.ref_msb_undo.argtypes = [C.c_uint32, u32p, u32p]
        L.ref_generate.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_double, C.c_uint32, u32p, C.c_size_t]
    return _cache["r"]


def ref_generate(kind, n, seed=0, rng=1 << 20, q=1.0, sub=0):
    """generate_inputs.cpp:70-117 through the reference shim. kind: 'zipf'|'geom'|'uniform'."""
    k = {"zipf": 0, "geom": 1, "uniform": 2}[kind]
    out = np.zeros(n, dtype=np.uint32)
    rc = reference().lib.ref_generate(k, seed, rng, q, sub, ptr(out, u32p), n)
    assert rc == 0
    return out


def mask_prelude_padding(blob, interp_end):
    """The reference's bit_stream leaves the unused high bits of its last 32-bit word
    uninitialised (bits.hpp:148, buf[3] is never zeroed).  Comparisons of preludes go
    through this: it returns a copy with the last interp word dropped.  `interp_end` is
    the byte offset just past the interp words."""
    b = np.array(blob, dtype=np.uint8, copy=True)
    b[interp_end - 4 : interp_end] = 0
    return b
