"""Drop-in check of the SONIC builder: `sonic-b200 --make-sonic` (conga_b200/host/sonic_host.cpp: libsonic's
sonic_build with its GC-profile loop replaced by cg_gc_profile at link time) against the UNMODIFIED builder
(oracle/_ref/sonic) on the same FASTA, gap / duplication BEDs and RepeatMasker file.  The containers must
hold the same bytes once inflated, the 8-byte build time (sonic/sonic.c:55) aside."""
import gzip
import os
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

import cli_inputs
import gc_cases

pytestmark = pytest.mark.gpu
HOST = os.path.join(cli_inputs.ROOT, "conga_b200", "host", "sonic-b200")
REF_SONIC = os.path.join(cli_inputs.REF, "sonic")


def write_fai(fa, text):
    with open(fa + ".fai", "w") as f:
        off = 0
        for block in text.split(b">")[1:]:
            head, _, body = block.partition(b"\n")
            lines = body.split(b"\n")
            w = len(lines[0]) if lines and lines[0] else 1
            off += 1 + len(head) + 1
            f.write("%s\t%d\t%d\t%d\t%d\n" % (head.split()[0].decode(), sum(len(l) for l in lines), off, w, w + 1))
            off += len(body)


def build_both(tmp, text, gaps=b"", dups=b"", reps=None, fai_text=None):
    fa = os.path.join(tmp, "ref.fa")
    open(fa, "wb").write(text)
    write_fai(fa, text if fai_text is None else fai_text)
    open(os.path.join(tmp, "gaps.bed"), "wb").write(gaps)
    open(os.path.join(tmp, "dups.bed"), "wb").write(dups)
    head = (b"   SW   perc perc perc  query     position in query    matching  repeat       position in repeat\n"
            b"score   div. del. ins.  sequence  begin end   (left)   repeat    class/family begin end (left) ID\n\n")
    open(os.path.join(tmp, "rm.out"), "wb").write(head + (reps or b""))
    out = {}
    for tag, exe in (("ref", REF_SONIC), ("b200", HOST)):
        sonic = os.path.join(tmp, tag + ".sonic")
        p = subprocess.run([exe, "--ref", fa, "--gaps", os.path.join(tmp, "gaps.bed"), "--dups", os.path.join(tmp, "dups.bed"),
                            "--reps", os.path.join(tmp, "rm.out"), "--info", "parity", "--make-sonic", sonic],
                           cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
        assert p.returncode == 0, "%s failed:\n%s" % (tag, p.stderr.decode(errors="replace")[-2000:])
        d = bytearray(gzip.open(sonic, "rb").read())
        d[4:12] = b"\0" * 8                                  # build time
        out[tag] = bytes(d)
    return out


@pytest.fixture(scope="module")
def tools():
    if not os.path.exists(REF_SONIC):
        pytest.fail("oracle/_ref/sonic missing: run `make -C oracle ref` where /root/reference exists")
    if not os.path.exists(HOST):
        pytest.fail("conga_b200/host/sonic-b200 missing: run `make -C conga_b200/host`")


def test_golden_fasta_texts(tools):
    g = np.load(gc_cases.GOLDEN)
    for case in sorted({k.split("/")[0] for k in g.files}):
        tmp = tempfile.mkdtemp(prefix="sonic_")
        try:
            out = build_both(tmp, g[case + "/fasta"].tobytes())
            assert out["ref"] == out["b200"], case
            assert len(out["ref"]) > 100
        finally:
            shutil.rmtree(tmp, ignore_errors=True)


def test_annotations_lower_case_crlf_and_unlisted_record(tools):
    """Gap / duplication / repeat sections present (written by the builder's own code on both sides), CR LF
    line ends, soft-masked bases, a description on the '>' line, and a record the .fai does not list."""
    rng = np.random.default_rng(77)
    alpha = np.frombuffer(b"ACGTacgtNn", dtype=np.uint8)

    def rec(name, L, width, eol=b"\n"):
        s = alpha[rng.integers(0, len(alpha), size=L)].tobytes()
        return b">" + name + b" desc text" + eol + eol.join(s[i:i + width] for i in range(0, L, width)) + eol
    listed = rec(b"1", 250_123, 60) + rec(b"2", 1_400_077, 70, b"\r\n") + rec(b"MT", 16_569, 61)
    text = listed + rec(b"unplaced", 5_000, 60)
    gaps = b"1\t1000\t2000\n2\t5000\t9000\n"
    dups = b"1\t30000\t42000\n2\t100000\t130000\n"
    reps = (b"  900  10.0  0.5  0.5  1  5000  6500  (243623)  +  REP1  Satellite/centr  1  1500  (0)  1\n"
            b"  800  12.0  0.5  0.5  2  70000  70300  (1329777)  C  AluY  SINE/Alu  (10)  300  1  2\n")
    tmp = tempfile.mkdtemp(prefix="sonic_")
    try:
        out = build_both(tmp, text, gaps, dups, reps, fai_text=listed)
        assert out["ref"] == out["b200"]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def test_chromosome_sized_record(tools):
    """60 Mb in two records: the multi-threaded white-space removal and a grid-stride launch."""
    rng = np.random.default_rng(78)
    alpha = np.frombuffer(b"ACGTacgtNN", dtype=np.uint8)
    parts = []
    for name, L in ((b"1", 45_000_017), (b"2", 15_000_000)):
        s = alpha[rng.integers(0, len(alpha), size=L)]
        full = (L // 60) * 60
        body = np.concatenate([s[:full].reshape(-1, 60), np.full((full // 60, 1), 10, dtype=np.uint8)], axis=1).tobytes()
        parts.append(b">" + name + b"\n" + body + s[full:].tobytes() + b"\n")
    tmp = tempfile.mkdtemp(prefix="sonic_")
    try:
        out = build_both(tmp, b"".join(parts))
        assert out["ref"] == out["b200"]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
