"""orc_gc_profile (oracle/conga_oracle.c) pinned against the UNMODIFIED reference SONIC builder: the GC bytes
`sonic --make-sonic` stored for the committed FASTA texts (tests/golden/gc/ref_gc.npz, made by
tests/golden/make_golden_gc.py)."""
import numpy as np

import gc_cases


def test_oracle_gc_profile_matches_reference_builder(oracle):
    n = 0
    for case, name, body, bases, want in gc_cases.golden_cases():
        assert len(want) == (len(bases) + 1) // 100, (case, name)
        got = oracle.gc_profile(body)                       # with the line feeds, as the builder reads it
        assert np.array_equal(got, want), (case, name)
        assert np.array_equal(oracle.gc_profile(bases), want), (case, name)
        assert np.array_equal(gc_cases.window_sums(bases), want), (case, name)
        n += 1
    assert n >= 18


def test_oracle_gc_profile_small_records(oracle):
    assert len(oracle.gc_profile(b"")) == 0
    assert len(oracle.gc_profile(b"G" * 98)) == 0
    assert oracle.gc_profile(b"G" * 99).tolist() == [99]
    assert oracle.gc_profile(b"g" * 99 + b"C" * 100).tolist() == [99, 100]
    assert oracle.gc_profile(b"N" * 99 + b"c" + b"A" * 99).tolist() == [0, 1]
