"""GPU, collected last: rcq_dust_mask_static — the DUST seeding mask computed with the library's static device
scratch (no allocation inside the call) — against the existing entry points (rcq_sample_create +
rcq_sample_seed_mask), bit for bit, on several calls in a row that share the scratch array."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LET = np.frombuffer(b"ACGT", dtype=np.uint8)


def _transcripts(seed, n_tx, low_complexity):
    """Random transcripts; a share of them carries poly-A tails, (CA)n / (CAG)n stretches and a few N's."""
    rng = np.random.default_rng(seed)
    lens = [int(x) for x in rng.integers(40, 2500, size=n_tx)]
    seqs = []
    for k, n in enumerate(lens):
        a = LET[rng.integers(0, 4, size=n)].copy()
        if low_complexity and k % 3 == 0 and n > 120:
            t = int(rng.integers(20, 90))
            a[n - t:] = ord("A")
        if low_complexity and k % 4 == 1 and n > 300:
            unit = (b"CA", b"CAG", b"AT", b"GGA")[(k // 4) % 4]
            m = int(rng.integers(30, 150))
            p = int(rng.integers(0, n - m))
            a[p:p + m] = np.frombuffer((unit * (m // len(unit) + 1))[:m], dtype=np.uint8)
        if k % 7 == 0 and n > 50:
            a[rng.integers(0, n, size=2)] = ord("N")
        seqs.append(a)
    ascii_ = np.concatenate(seqs)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    return ascii_, offsets


def _reference_mask(ctx, seed, n_tx, low_complexity=True):
    """The existing path: build a sample, read back its packed arena and its seeding mask."""
    from rna_clique_b200 import _native
    ascii_, offsets = _transcripts(seed, n_tx, low_complexity)
    ids = np.arange(n_tx, dtype=np.int32)
    s = _native.Sample(ctx, ascii_, offsets, ids, np.ones(n_tx, dtype=np.int32))
    d = s.download()
    smask, n_dust = s.seed_mask()
    arena = s.arena
    s.close()
    return d["fwd"], d["mask"], smask, n_dust, arena


def _static_mask(fwd, mask, arena, stream=None, want_count=True):
    """The new entry point through the C ABI: the caller owns inputs and output (torch tensors stand in for any
    device buffers), the library brings its own scratch."""
    import torch
    from rna_clique_b200 import _native
    dev = torch.device("cuda", 0)
    fwd_d = torch.from_numpy(fwd.view(np.int32)).to(dev)
    mask_d = torch.from_numpy(mask.view(np.int32)).to(dev)
    out_d = torch.full((arena // 32,), 0x5A5A5A5A, dtype=torch.int32, device=dev)   # stale contents must not survive
    torch.cuda.synchronize(dev)
    n = C.c_int64(-1)
    rc = _native.lib().rcq_dust_mask_static(0, fwd_d.data_ptr(), mask_d.data_ptr(), arena, out_d.data_ptr(), stream,
                                            C.byref(n) if want_count else None)
    assert rc == 0, _native.lib().rcq_last_error().decode()
    torch.cuda.synchronize(dev)
    return out_d.cpu().numpy().view(np.uint32), int(n.value)


def test_static_scratch_dust_mask_equals_the_sample_build_on_calls_in_a_row():
    from rna_clique_b200 import _native
    ctx = _native.Context(0)
    try:
        big = _reference_mask(ctx, seed=5, n_tx=900)
        small = _reference_mask(ctx, seed=6, n_tx=60)
        clean = _reference_mask(ctx, seed=7, n_tx=200, low_complexity=False)
    finally:
        ctx.close()
    assert big[3] > 5000 and small[3] > 100                    # DUST has something to do
    assert big[4] > small[4]                                    # the second call leaves a stale tail in the scratch
    # two calls in a row, then more: every call shares the one scratch array
    for fwd, mask, want, n_dust, arena in (big, big, small, clean, big):
        got, n = _static_mask(fwd, mask, arena)
        assert np.array_equal(got, want)
        assert n == n_dust
    # asynchronous form: no count asked for, the caller synchronises
    got, _ = _static_mask(small[0], small[1], small[4], want_count=False)
    assert np.array_equal(got, small[2])


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    import torch
    from rna_clique_b200 import _native
    L = _native.lib()
    limit = 1 << 24
    buf = torch.zeros(16, dtype=torch.int32, device="cuda:0")   # never touched: the size check comes first
    rc = L.rcq_dust_mask_static(0, buf.data_ptr(), buf.data_ptr(), limit + 64, buf.data_ptr(), None, None)
    assert rc == -3                                             # RCQ_ERR_CAPACITY
    msg = L.rcq_last_error().decode()
    assert str(limit) in msg and "static scratch" in msg
    assert L.rcq_dust_mask_static(0, buf.data_ptr(), buf.data_ptr(), 100, buf.data_ptr(), None, None) == -2   # not a multiple of 64
    assert L.rcq_dust_mask_static(0, None, buf.data_ptr(), 64, buf.data_ptr(), None, None) == -2
