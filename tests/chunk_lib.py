"""H2O chunk byte arrays for the ingest tests, built and decoded by the CPU oracle (oracle/h2o_chunk_oracle.hpp).
TEST INFRASTRUCTURE ONLY."""
import ctypes as C

import numpy as np

import oracle_lib as O

(C0D, C0L, C1, C1N, C1S, C2, C2S, C4, C4S, C4F, C8, C8D, CBS, CUD, CXI, CXF) = range(16)
NAMES = ["C0D", "C0L", "C1", "C1N", "C1S", "C2", "C2S", "C4", "C4S", "C4F", "C8", "C8D", "CBS", "CUD", "CXI", "CXF"]
_ready = False


def _lib():
    global _ready
    L = O.lib()
    if not _ready:
        L.orc_chunk_rows.restype = C.c_int64
        L.orc_chunk_rows.argtypes = [C.c_int, C.c_void_p, C.c_int64]
        L.orc_chunk_decode.restype = None
        L.orc_chunk_decode.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        L.orc_chunk_enc_int.restype = C.c_int64
        L.orc_chunk_enc_int.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_chunk_enc_scaled.restype = C.c_int64
        L.orc_chunk_enc_scaled.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_chunk_enc_c0.restype = C.c_int64
        L.orc_chunk_enc_c0.argtypes = [C.c_int, C.c_double, C.c_int64, C.c_int32, C.c_void_p, C.c_int64]
        L.orc_chunk_enc_cbs.restype = C.c_int64
        L.orc_chunk_enc_cbs.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64]
        L.orc_chunk_enc_cud.restype = C.c_int64
        L.orc_chunk_enc_cud.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_chunk_enc_cxi.restype = C.c_int64
        L.orc_chunk_enc_cxi.argtypes = [C.c_int32, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_int64]
        L.orc_chunk_enc_cxf.restype = C.c_int64
        L.orc_chunk_enc_cxf.argtypes = [C.c_int32, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_pow10.restype = C.c_double
        L.orc_pow10.argtypes = [C.c_int64, C.c_int]
        _ready = True
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _run(fn, cap, *args):
    out = np.zeros(max(int(cap), 16), dtype=np.uint8)
    n = fn(*args, _p(out), out.size)
    assert n >= 0, "encoder output larger than the buffer"
    return out[:n].copy()


def rows(codec, mem):
    mem = np.ascontiguousarray(mem, dtype=np.uint8)
    return int(_lib().orc_chunk_rows(codec, _p(mem), mem.size))


def decode(codec, mem, n=None):
    """Chunk.getDoubles of the oracle: float64[n], NaN == NA."""
    mem = np.ascontiguousarray(mem, dtype=np.uint8)
    if n is None:
        n = rows(codec, mem)
    out = np.zeros(n, dtype=np.float64)
    _lib().orc_chunk_decode(codec, _p(mem), mem.size, n, _p(out))
    return out


def enc_int(log, vals, na=None):
    v = np.ascontiguousarray(vals, dtype=np.int64)
    m = np.zeros(v.size, np.uint8) if na is None else np.ascontiguousarray(na, dtype=np.uint8)
    return _run(_lib().orc_chunk_enc_int, v.size << log, log, v.size, _p(v), _p(m))


def enc_scaled(log, man, exp, na=None):
    a = np.ascontiguousarray(man, dtype=np.int64)
    x = np.ascontiguousarray(exp, dtype=np.int32)
    m = np.zeros(a.size, np.uint8) if na is None else np.ascontiguousarray(na, dtype=np.uint8)
    return _run(_lib().orc_chunk_enc_scaled, 16 + (a.size << log), log, a.size, _p(a), _p(x), _p(m))


def enc_c0(con, n, is_long=False):
    return _run(_lib().orc_chunk_enc_c0, 12, 1 if is_long else 0, float(con), int(con) if is_long else 0, n)


def enc_cbs(vals, na=None, bpv=None):
    v = np.ascontiguousarray(vals, dtype=np.uint8)
    m = np.zeros(v.size, np.uint8) if na is None else np.ascontiguousarray(na, dtype=np.uint8)
    if bpv is None:
        bpv = 2 if m.any() else 1
    return _run(_lib().orc_chunk_enc_cbs, 3 + v.size, v.size, _p(v), _p(m), bpv)


def enc_cud(vals):
    v = np.ascontiguousarray(vals, dtype=np.float64)
    return _run(_lib().orc_chunk_enc_cud, 8 + 8 * 256 + v.size, v.size, _p(v))


def enc_cxi(n, id_sz, val_sz, na_sparse, ids, vals, na=None):
    i = np.ascontiguousarray(ids, dtype=np.int32)
    v = np.ascontiguousarray(vals, dtype=np.int64)
    m = np.zeros(i.size, np.uint8) if na is None else np.ascontiguousarray(na, dtype=np.uint8)
    return _run(_lib().orc_chunk_enc_cxi, 8 + i.size * (id_sz + val_sz), n, id_sz, val_sz, int(na_sparse), i.size, _p(i), _p(v), _p(m))


def enc_cxf(n, valsz, na_sparse, ids, vals):
    i = np.ascontiguousarray(ids, dtype=np.int32)
    v = np.ascontiguousarray(vals, dtype=np.float64)
    return _run(_lib().orc_chunk_enc_cxf, 8 + i.size * (4 + valsz), n, valsz, int(na_sparse), i.size, _p(i), _p(v))


def pow10(m, e):
    return float(_lib().orc_pow10(int(m), int(e)))
