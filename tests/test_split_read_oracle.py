"""The oracle's split-read restatement (oracle/conga_oracle.c: orc_sr_rows, orc_sr_count) and the Python
restatement of the host-side gate (tests/split_ref.py), pinned against the UNMODIFIED reference: the per-CNV
rp / border_rp values its --wrap dump produced with --rp (tests/golden/rp/ref_rp.npz, made by
tests/golden/make_golden_rp.py)."""
import os

import numpy as np

import split_ref
from conga_b200.abi import CNV_DT

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rp", "ref_rp.npz")


def chain(oracle, g, k):
    L = int(g["chrom_lens"][k])
    mq, min_len = int(g["mq_threshold"]), int(g["min_read_length"])
    rep = g["c%d_repeats" % k]
    halves, bases = split_ref.halves_of(g["c%d_reads" % k], g["c%d_seqqual" % k], rep, mq, min_len)
    rows, n_map = oracle.sr_rows(g["c%d_ref" % k], halves, bases, L, mq)
    rows = split_ref.filter_and_widen(rows, rep)
    dels, dups = np.zeros(len(g["c%d_del" % k]), dtype=CNV_DT), np.zeros(len(g["c%d_dup" % k]), dtype=CNV_DT)
    for a, src in ((dels, g["c%d_del" % k]), (dups, g["c%d_dup" % k])):
        a["start"], a["end"] = src["start"], src["end"]
    return halves, rows, n_map, oracle.sr_count(rows, dels, dups)


def test_oracle_split_read_chain_matches_reference_dump(oracle):
    g = np.load(GOLDEN)
    seen_rp = seen_border = 0
    for k in range(len(g["chrom_names"])):
        halves, rows, n_map, (border_rp, rp) = chain(oracle, g, k)
        assert len(halves) > 1000 and n_map >= len(halves) // 2
        assert np.array_equal(border_rp, g["c%d_del" % k]["border_rp"])
        assert np.array_equal(rp, g["c%d_dup" % k]["rp"])
        seen_rp += int(rp.sum())
        seen_border += int(border_rp.sum())
    assert seen_rp > 50 and seen_border > 50


def test_reference_loaded_length_quirk_is_in_the_fixture():
    g = np.load(GOLDEN)
    # chromosome 0 is 180 700 long: readReferenceSeq stops at 180 000 (common.c:434)
    assert int(g["chrom_lens"][0]) == 180_700 and len(g["c0_ref"]) == 180_000
    assert len(g["c1_ref"]) == int(g["chrom_lens"][1])
