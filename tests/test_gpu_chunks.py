"""Frame ingest from H2O chunks: h2ot_load_chunk decodes Chunk._mem on the device; every codec must give
exactly what Chunk.atd_impl gives (oracle/h2o_chunk_oracle.hpp, pinned in tests/test_oracle_chunks.py), and a model built
from chunks must be the model built from the decoded doubles."""
import numpy as np
import pytest

import chunk_lib as K
import h2ot_lib

pytestmark = pytest.mark.gpu
NaN = float("nan")


def _chunks_for(codec, n, rng):
    """One chunk of n rows in the given codec with random content: (mem bytes, decoded doubles by the oracle)."""
    na = rng.random(n) < 0.05
    if codec == K.C0D:
        mem = K.enc_c0(rng.choice([2.5, NaN, -1e-3]), n)
    elif codec == K.C0L:
        mem = K.enc_c0(int(rng.integers(-10 ** 12, 10 ** 12)), n, is_long=True)
    elif codec == K.C1:
        mem = K.enc_int(0, rng.integers(0, 255, n), na)
    elif codec == K.C1N:
        mem = K.enc_int(0, rng.integers(0, 256, n))
    elif codec == K.C2:
        mem = K.enc_int(1, rng.integers(-32767, 32768, n), na)
    elif codec == K.C4:
        mem = K.enc_int(2, rng.integers(-2 ** 31 + 1, 2 ** 31, n), na)
    elif codec == K.C8:
        mem = K.enc_int(3, rng.integers(-2 ** 62, 2 ** 62, n), na)
    elif codec == K.C1S:
        e = int(rng.choice([-3, -1, 2]))
        mem = K.enc_scaled(0, int(rng.integers(-5000, 5000)) + rng.integers(0, 254, n), np.full(n, e), na)
    elif codec == K.C2S:
        e = int(rng.choice([-4, -2, 1]))
        mem = K.enc_scaled(1, int(rng.integers(-10 ** 6, 10 ** 6)) + rng.integers(0, 65000, n), np.full(n, e), na)
    elif codec == K.C4S:
        e = int(rng.choice([-9, -6, -2]))
        mem = K.enc_scaled(2, int(rng.integers(-10 ** 9, 10 ** 9)) + rng.integers(0, 2 ** 32 - 10, n), np.full(n, e), na)
    elif codec == K.C4F:
        v = rng.normal(size=n).astype(np.float32)
        v[na] = np.float32(NaN)
        mem = np.frombuffer(v.tobytes(), dtype=np.uint8)
    elif codec == K.C8D:
        v = rng.lognormal(size=n)
        v[na] = NaN
        mem = np.frombuffer(v.tobytes(), dtype=np.uint8)
    elif codec == K.CBS:
        mem = K.enc_cbs(rng.integers(0, 2, n), na if rng.random() < 0.5 else None)
    elif codec == K.CUD:
        pool = np.concatenate([rng.normal(size=40), [NaN, 0.0, 1e300]])
        mem = K.enc_cud(pool[rng.integers(0, pool.size, n)])
    elif codec == K.CXI:
        id_sz, val_sz = [(2, 0), (4, 0), (2, 2), (4, 4), (4, 8)][int(rng.integers(0, 5))]
        if id_sz == 2 and n > 65535:
            id_sz, val_sz = 4, 4
        nnz = max(1, n // 20)
        ids = np.sort(rng.choice(n, nnz, replace=False))
        hi = {0: 2, 2: 2 ** 15, 4: 2 ** 31, 8: 2 ** 62}[val_sz]
        vals = rng.integers(-hi + 1, hi, nnz) if val_sz else np.ones(nnz, np.int64)
        mem = K.enc_cxi(n, id_sz, val_sz, bool(rng.integers(0, 2)), ids, vals, (rng.random(nnz) < 0.1) if val_sz else None)
    elif codec == K.CXF:
        nnz = max(1, n // 20)
        ids = np.sort(rng.choice(n, nnz, replace=False))
        vals = rng.normal(size=nnz)
        vals[::13] = NaN
        mem = K.enc_cxf(n, int(rng.choice([4, 8])), bool(rng.integers(0, 2)), ids, vals)
    else:
        raise AssertionError(codec)
    assert K.rows(codec, mem) == n, K.NAMES[codec]
    return mem, K.decode(codec, mem, n)


def test_every_codec_decodes_like_atd():
    """One column per codec plus mixed-codec columns (every Chunk of a Vec picks its own codec), three chunks each."""
    P = h2ot_lib.pkg()
    rng = np.random.default_rng(21)
    sizes = [40000, 1, 25537]
    n = sum(sizes)
    plans = [[c] * 3 for c in range(16)] + [[K.C8D, K.CUD, K.C4S], [K.C1, K.C2S, K.CXI], [K.CBS, K.C0L, K.C1N], [K.CXF, K.C0D, K.C4F]]
    e = P.Engine(nrows=n, ncols=len(plans), family="gaussian")
    want = []
    for col, plan in enumerate(plans):
        row0, parts = 0, []
        for codec, m in zip(plan, sizes):
            mem, dec = _chunks_for(codec, m, rng)
            e.load_chunk(col, codec, mem, row0, m)
            parts.append(dec)
            row0 += m
        want.append(np.concatenate(parts))
    ymem, ydec = _chunks_for(K.C4F, n, rng)
    e.load_chunk(-1, K.C4F, ymem, 0, n)
    e.finalize_frame()
    for col, w in enumerate(want):
        got = e.get_column(col)
        assert np.array_equal(got, w, equal_nan=True), (col, [K.NAMES[c] for c in plans[col]], int(np.sum(~((got == w) | (np.isnan(got) & np.isnan(w))))))
        info = e.column_info(col)
        assert info.nacnt == int(np.isnan(w).sum())
        if np.any(~np.isnan(w)):
            assert info.min == np.nanmin(w) and info.max == np.nanmax(w)
    e.close()


def test_malformed_chunks_are_rejected():
    P = h2ot_lib.pkg()
    e = P.Engine(nrows=100, ncols=1, family="gaussian")
    good = K.enc_int(1, np.arange(100))
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, K.C2, good, 0, 99)            # the chunk holds 100 rows
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, K.C2, good[:-1], 0, 100)      # odd byte count
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, K.C2, good, 1, 100)           # past the end of the column
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, 99, good, 0, 100)             # unknown codec
    bad = K.enc_cxi(100, 4, 4, False, [5, 100], [1, 2])
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, K.CXI, bad, 0, 100)           # row id 100 out of range
    odd = K.enc_cxi(100, 4, 2, False, [5], [1])
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, K.CXI, odd, 0, 100)           # (4-byte id, 2-byte value) is not a layout CXIChunk reads
    s = K.enc_scaled(1, np.arange(100), np.full(100, -2))
    s[12] = 7                                         # size log that contradicts the codec
    with pytest.raises(P.H2OTError):
        e.load_chunk(0, K.C2S, s, 0, 100)
    e.load_chunk(0, K.C2, good, 0, 100)               # the engine is still usable
    e.load_response(np.zeros(100))
    e.finalize_frame()
    assert np.array_equal(e.get_column(0), np.arange(100.0))
    e.close()


def test_model_from_chunks_equals_model_from_doubles():
    """GBM on a frame pushed chunk by chunk (mixed codecs, categorical via C1/C2, response as CBS, weights as C1N) ==
    GBM on the same frame pushed as doubles: identical trees and predictions."""
    P = h2ot_lib.pkg()
    rng = np.random.default_rng(22)
    sizes = [30000, 20000, 10001]
    n = sum(sizes)
    plans = [[K.C8D] * 3, [K.C4F, K.C8D, K.C4F], [K.C2S] * 3, [K.C1, K.C1, K.C1N], [K.C2] * 3, [K.CXI, K.C0L, K.C1], [K.CUD] * 3, [K.C4S] * 3]
    cols_chunks, cols_dec = [], []
    for plan in plans:
        ch, parts = [], []
        for codec, m in zip(plan, sizes):
            mem, dec = _chunks_for(codec, m, rng)
            ch.append((codec, mem, m))
            parts.append(dec)
        cols_chunks.append(ch)
        cols_dec.append(np.concatenate(parts))
    X = np.stack(cols_dec, axis=1)
    z = np.nan_to_num(X[:, 0]) * 0.5 - np.nan_to_num(X[:, 1]) + (np.nan_to_num(X[:, 3]) % 5) * 0.3
    y = (z + rng.normal(size=n) > np.median(z)).astype(np.float64)
    w = rng.integers(1, 4, n).astype(np.float64)
    ych, wch, r0 = [], [], 0
    for m in sizes:
        ych.append((K.CBS, K.enc_cbs(y[r0:r0 + m].astype(np.uint8)), m))
        wch.append((K.C1N, K.enc_int(0, w[r0:r0 + m].astype(np.int64)), m))
        r0 += m
    cat = {3: 256}
    kw = dict(ntrees=5, max_depth=5, nbins=20, distribution="bernoulli", seed=1234)
    a = P.GBM(**kw).load([X[:, c] for c in range(X.shape[1])], y, weights=w, categorical=cat).train()
    b = P.GBM(**kw).load_chunks(n, cols_chunks, ych, weights=wch, categorical=cat).train()
    assert a.ntrees == b.ntrees == 5
    for t in range(5):
        na, nb = a.tree(t)[0], b.tree(t)[0]
        assert len(na) == len(nb)
        for x, z2 in zip(na, nb):
            for f, _ in P.TreeNode._fields_:
                if f == "bs":
                    assert bytes(x.bs[:x.bs_nbytes]) == bytes(z2.bs[:z2.bs_nbytes])
                else:
                    u, v = getattr(x, f), getattr(z2, f)
                    assert u == v or (u != u and v != v), (t, f, u, v)
    assert np.array_equal(a.predictions(), b.predictions())
    a.close()
    b.close()
