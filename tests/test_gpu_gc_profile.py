"""k_gc_profile (conga_b200/csrc/gc_profile.cu) through cg_gc_profile against the reference builder's golden
bytes, the oracle and, at chromosome size, the NumPy window rule — HOST and DEVICE pointer modes."""
import numpy as np
import pytest
import torch

import gc_cases
from conga_b200 import CG_PTR_DEVICE, CG_PTR_HOST

pytestmark = pytest.mark.gpu


def test_golden_records_host_mode(ctx, oracle):
    n = 0
    for case, name, body, bases, want in gc_cases.golden_cases():
        got = ctx.gc_profile(bases)
        assert got.dtype == np.int8 and np.array_equal(got, want), (case, name)
        n += 1
    assert n >= 18


def test_small_and_empty_records(ctx):
    assert len(ctx.gc_profile(b"")) == 0
    assert len(ctx.gc_profile(b"G" * 98)) == 0
    assert ctx.gc_profile(b"G" * 99).tolist() == [99]
    assert ctx.gc_profile(b"g" * 99 + b"C" * 100).tolist() == [99, 100]
    assert ctx.gc_profile(b"N" * 99 + b"c" + b"A" * 99 + b"GGG").tolist() == [0, 1]


def random_bases(rng, n):
    alpha = np.frombuffer(b"ACGTacgtNnRYSWKM", dtype=np.uint8)
    return alpha[rng.integers(0, len(alpha), size=n)]


@pytest.mark.parametrize("n", [12_799, 12_800, 12_899, 1_000_003, 30_000_000])
def test_random_records_against_oracle(ctx, oracle, n):
    rng = np.random.default_rng(n)
    b = random_bases(rng, n)
    got = ctx.gc_profile(b)
    assert np.array_equal(got, oracle.gc_profile(b.tobytes()))
    assert np.array_equal(got, gc_cases.window_sums(b.tobytes()))


def test_device_mode_unaligned_pointer_and_chromosome_size(ctx):
    """DEVICE pointers, a base pointer that is not 16-byte aligned, chr1-sized input (249 M bases)."""
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev); g.manual_seed(5)
    n = 249_250_621
    codes = torch.randint(0, 8, (n + 3,), device=dev, dtype=torch.uint8, generator=g)
    alpha = torch.tensor(list(b"ACGTgcNn"), dtype=torch.uint8, device=dev)
    buf = alpha[codes.long()] if n < 1 << 20 else torch.take(alpha, codes.to(torch.int64))
    del codes
    bases = buf[3:]                                          # address = 256-aligned + 3
    out = torch.full(((n + 1) // 100,), -7, dtype=torch.int8, device=dev)
    ctx.set_pointer_mode(CG_PTR_DEVICE)
    try:
        ctx.gc_profile(bases, n_bases=n, out=out)
        ctx.sync()
    finally:
        ctx.set_pointer_mode(CG_PTR_HOST)
    is_gc = ((bases == ord("G")) | (bases == ord("C")) | (bases == ord("g")) | (bases == ord("c")))
    del buf
    cs = torch.cumsum(is_gc.to(torch.int32), 0, dtype=torch.int32)
    cs = torch.cat([torch.zeros(1, dtype=torch.int32, device=dev), cs])
    w = torch.arange((n + 1) // 100, device=dev, dtype=torch.int64)
    want = (cs[100 * w + 99] - cs[torch.clamp(100 * w - 1, min=0)]).to(torch.int8)
    assert torch.equal(out, want)
