"""Pins the chunk-codec oracle (oracle/h2o_chunk_oracle.hpp) with the vectors of the reference's own unit tests
(h2o-core/src/test/java/water/fvec/*ChunkTest.java): writer restatement -> reader restatement must give the values
those tests assert; for the scaled codecs that is Double.parseDouble("<m>e<x>") == Python's float("<m>e<x>")."""
import struct

import numpy as np
import pytest

import chunk_lib as K

NaN = float("nan")


def _eq(a, b):
    return np.array_equal(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), equal_nan=True)


@pytest.mark.parametrize("lead_na", [False, True])
def test_c1_c2_c4_c8_vectors(lead_na):
    """C1ChunkTest:18, C2ChunkTest:18, C4ChunkTest:17 (+ a C8 analogue): values, NA first / last."""
    for codec, log, vals in ((K.C1, 0, [0, 1, 3, 254]), (K.C2, 1, [-32767, 0, 32767]), (K.C4, 2, [-2147483647, 0, 2147483647]),
                             (K.C8, 3, [-(2 ** 63) + 1, 0, 2 ** 62])):
        v = ([0] if lead_na else []) + vals + [0]
        na = ([1] if lead_na else []) + [0] * len(vals) + [1]
        mem = K.enc_int(log, v, na)
        assert mem.size == len(v) << log and K.rows(codec, mem) == len(v)
        want = ([NaN] if lead_na else []) + [float(x) for x in vals] + [NaN]
        assert _eq(K.decode(codec, mem), want)
    mem = K.enc_int(0, [0, 7, 255])
    assert _eq(K.decode(K.C1N, mem), [0, 7, 255])  # C1NChunk has no NA: 0xFF is 255


def test_scaled_inflate_vectors():
    """C1SChunkTest:25-26 and :85-86, C2SChunkTest:25-26, C4SChunkTest:19-20: mantissa / exponent pairs."""
    cases = [(K.C1S, 0, [0, 2, 3, 254], [1, -1, -1, -2]), (K.C1S, 0, [-1228, -997, -9740], [-4, -4, -5]),
             (K.C2S, 1, [-32767, 34, 0, 32767], [-3, -2, 1, -3]),
             (K.C4S, 2, [-2147483647, 0, 215, 188001238, 2147483647], [-9, 1, -6, -8, -9])]
    for codec, log, man, exp in cases:
        for lead_na in (False, True):
            m = ([0] if lead_na else []) + man + [0]
            x = ([0] if lead_na else []) + exp + [0]
            na = ([1] if lead_na else []) + [0] * len(man) + [1]
            mem = K.enc_scaled(log, m, x, na)
            got = K.decode(codec, mem)
            want = ([NaN] if lead_na else []) + [K.pow10(a, b) for a, b in zip(man, exp)] + [NaN]
            assert _eq(got, want), (K.NAMES[codec], got, want)


def test_c1s_precision_known_answers():
    """C1SChunkTest.test_precision (:175-198): 255 consecutive mantissas from a bias, every exponent; the chunk must give
    exactly Double.parseDouble((i + bias) + "e" + exponent) -- the correctly rounded decimal, as Python's float() parses it."""
    for exponent in (-16, -8, -6, -4, -2, -1, 0, 1, 2, 4, 6, 8, 16):
        for bias in (-1234567, -12345, -1234, -1, 0, 1, 1234, 12345, 1234567):
            if exponent == 0 and abs(bias) <= 1:
                continue
            man = [bias + i for i in range(255)]
            mem = K.enc_scaled(0, man, [exponent] * 255)
            assert struct.unpack_from("<q", mem.tobytes(), 0)[0] == bias
            got = K.decode(K.C1S, mem)
            want = [float("%de%d" % (m, exponent)) for m in man]
            assert _eq(got, want), (exponent, bias)


def test_c4s_precision_known_answers():
    """C4SChunkTest (:133-150): same protocol over a 4-byte range."""
    for exponent in (-16, -8, -6, -4, -2, -1):
        for bias in (-1234567, -12345, -1234, -1, 0, 1, 1234, 12345, 1234567):
            man = [bias + i * 16843009 for i in range(255)]
            mem = K.enc_scaled(2, man, [exponent] * 255)
            got = K.decode(K.C4S, mem)
            want = [float("%de%d" % (m, exponent)) for m in man]
            assert _eq(got, want), (exponent, bias)


def test_cbs_vectors():
    """CBSChunkTest.test1BPV / test2BPV (:120-160): expected bits per value, gap, compressed length and values."""
    NA = None
    cases = [([1, 0, 1, 1], 1, 4, 1), ([1, 0, 0, 0, 1, 1, 1, 0], 1, 0, 1), ([1, 0, 0, 0, 1, 1, 1, 0, 1], 1, 7, 2),
             ([0, NA, 1], 2, 2, 1), ([1, NA, 0, 1], 2, 0, 1), ([1, 0, NA, 1, 0, 0], 2, 4, 2),
             ([NA, NA, NA, 1, 0, NA, 1, NA], 2, 0, 2)]
    for vals, bpv, gap, clen in cases:
        v = [0 if x is None else x for x in vals]
        na = [1 if x is None else 0 for x in vals]
        mem = K.enc_cbs(v, na)
        assert (mem[1], mem[0], mem.size - 2) == (bpv, gap, clen), vals
        assert K.rows(K.CBS, mem) == len(vals)
        assert _eq(K.decode(K.CBS, mem), [NaN if x is None else float(x) for x in vals])


def test_constant_unique_and_sparse_layouts():
    """C0D / C0L (C0DChunkTest, C0LChunkTest), CUD (CUDChunkTest), CXI / CXF (NewChunk.bufS / bufD): round trips."""
    assert _eq(K.decode(K.C0D, K.enc_c0(3.25, 5)), [3.25] * 5)
    assert _eq(K.decode(K.C0D, K.enc_c0(NaN, 3)), [NaN] * 3)
    assert _eq(K.decode(K.C0L, K.enc_c0(-77, 4, is_long=True)), [-77.0] * 4)
    rng = np.random.default_rng(5)
    pool = np.array([0.1, -2.5e300, 3.0, NaN, -0.0, 1e-310])
    v = pool[rng.integers(0, pool.size, 1000)]
    mem = K.enc_cud(v)
    assert mem.size == 8 + 8 * 6 + 1000  # CUDChunk.computeByteSize
    got = K.decode(K.CUD, mem)
    assert np.array_equal(got.view(np.int64), v.view(np.int64))  # bit for bit, -0.0 and NaN included
    n = 5000
    for id_sz, val_sz in ((2, 0), (4, 0), (2, 2), (4, 4), (4, 8)):
        for na_sparse in (False, True):
            ids = np.sort(rng.choice(n, 300, replace=False)).astype(np.int32)
            hi = {0: 2, 2: 2 ** 15, 4: 2 ** 31, 8: 2 ** 62}[val_sz]
            vals = rng.integers(-hi + 1, hi, ids.size) if val_sz else np.ones(ids.size, np.int64)
            na = (rng.random(ids.size) < 0.1) if val_sz else np.zeros(ids.size, bool)
            mem = K.enc_cxi(n, id_sz, val_sz, na_sparse, ids, vals, na)
            want = np.full(n, NaN if na_sparse else 0.0)
            want[ids] = np.where(na, NaN, vals.astype(np.float64))
            assert _eq(K.decode(K.CXI, mem), want), (id_sz, val_sz, na_sparse)
    for valsz in (4, 8):
        ids = np.sort(rng.choice(n, 300, replace=False)).astype(np.int32)
        vals = rng.normal(size=ids.size)
        vals[::17] = NaN
        mem = K.enc_cxf(n, valsz, True, ids, vals)
        want = np.full(n, NaN)
        want[ids] = vals.astype(np.float32).astype(np.float64) if valsz == 4 else vals
        assert _eq(K.decode(K.CXF, mem), want)
