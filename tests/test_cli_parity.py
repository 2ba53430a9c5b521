"""Drop-in check at the command line: `conga-b200` (conga_b200/host, the four seams routed through
libconga_gpu.so) against the UNMODIFIED reference program (oracle/_ref/conga) on the same BAM, SONIC and
BED files.  <out>_svs.bed, <out>_dels.bed, <out>_dups.bed and the conga.log chromosome lines must be
byte-identical."""
import os
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

import cli_inputs

pytestmark = pytest.mark.gpu
ROOT = cli_inputs.ROOT
HOST = os.environ.get("CONGA_HOST_BIN") or os.path.join(ROOT, "conga_b200", "host", "conga-b200")   # CONGA_HOST_BIN: the sanitizer build (scripts/gpu_asan.sh)
REF_CONGA = os.path.join(cli_inputs.REF, "conga")


def run_both(tmp, paths, flags, env=None):
    outs = {}
    for tag, exe in (("ref", REF_CONGA), ("b200", HOST)):
        wd = os.path.join(tmp, tag)
        os.makedirs(os.path.join(wd, "o"), exist_ok=True)
        cmd = [exe, "-i", paths["bam"], "--ref", paths["ref"], "--sonic", paths["sonic"], "--out", os.path.join(wd, "o", "s")] + flags
        p = subprocess.run(cmd, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600,
                           env=dict(os.environ, **(env or {})) if tag == "b200" else None)
        assert p.returncode == 0, "%s failed:\n%s" % (tag, p.stderr.decode()[-2000:])
        files = {}
        for nm in ("s_svs.bed", "s_dels.bed", "s_dups.bed"):
            f = os.path.join(wd, "o", nm)
            files[nm] = open(f, "rb").read() if os.path.exists(f) else None
        log = open(os.path.join(wd, "conga.log"), "rb").read().split(b"\n")
        # line 0 is the creation date; the BAM/ref/SONIC/first/last lines and one line per chromosome follow
        files["log"] = b"\n".join(log[1:])
        outs[tag] = files
    return outs


def assert_same(outs, expect_rows=True):
    for nm in outs["ref"]:
        assert outs["ref"][nm] == outs["b200"][nm], "%s differs from the reference's" % nm
    if expect_rows:
        assert outs["ref"]["s_svs.bed"].count(b"\n") > 1, "the case genotypes nothing: not a test"


@pytest.fixture(scope="module")
def tools():
    if not cli_inputs.have_reference_binaries():
        pytest.fail("oracle/_ref binaries missing: run `make -C oracle ref` where /root/reference exists")
    if not os.path.exists(HOST):
        pytest.fail("conga_b200/host/conga-b200 missing: run `make -C conga_b200/host`")


def test_plain_dels_and_dups(tools):
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(101)
        paths = cli_inputs.build_inputs(tmp, [("1", 400_000), ("2", 250_000), ("X", 100_000), ("MT", 16_500)], rng)
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--dups", paths["dups"]])
        assert_same(outs)
        assert b"DEL" in outs["ref"]["s_svs.bed"] and b"X\t" not in outs["ref"]["s_dels.bed"]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def test_mappability_min_mapq_c_score(tools):
    """... and CONGA_STATS=<file>: the opt-in machine-readable summary leaves every output as it is."""
    import json
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(102)
        paths = cli_inputs.build_inputs(tmp, [("1", 300_000), ("chr2", 200_000)], rng, reads_per_mb=60000, with_map=True)
        stats = os.path.join(tmp, "stats.json")
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--dups", paths["dups"], "--mappability", paths["map"],
                                     "--min-mapq", "20", "--c-score", "0.3", "--min-sv-size", "1200"], env={"CONGA_STATS": stats})
        assert_same(outs)
        st = json.load(open(stats))
        assert st["chromosomes"] == 2 and st["positions"] == 500_000 and st["reads"] > 0 and st["cnvs"] > 0
        assert st["kernel_launches"]["depth_pass"] == 2 and st["kernel_ms"]["depth_pass"] > 0 and st["depth_pass_gb_per_s"] > 0
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def test_chromosome_range_missing_target_and_dels_only(tools):
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(103)
        chroms = [("1", 200_000), ("2", 150_000), ("3", 120_000), ("4", 100_000)]
        paths = cli_inputs.build_inputs(tmp, chroms, rng, bam_chroms={"1", "2", "4"})        # "3" is not in the BAM
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--first-chr", "1", "--last-chr", "3"])
        assert_same(outs)
        assert outs["ref"]["s_dups.bed"] is None and b"\n1\t" not in outs["ref"]["s_dels.bed"]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def test_int16_wrap_pileup_through_the_cli(tools):
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(104)
        paths = cli_inputs.build_inputs(tmp, [("1", 200_000)], rng, reads_per_mb=3000, pileup=True)
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--dups", paths["dups"]])
        assert_same(outs)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def test_two_gpus_write_the_same_bytes(tools):
    """CONGA_GPUS=2: chromosomes spread over two GPUs (LPT), NCCL all-reduce of the statistics table, outputs
    in SONIC order — still byte-identical to the single-threaded reference."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(105)
        chroms = [("1", 300_000), ("2", 260_000), ("3", 200_000), ("4", 150_000), ("5", 90_000), ("X", 50_000)]
        paths = cli_inputs.build_inputs(tmp, chroms, rng, with_map=True, bam_chroms={"1", "2", "3", "5", "X"})
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--dups", paths["dups"], "--mappability", paths["map"]],
                        env={"CONGA_GPUS": "2"})
        assert_same(outs)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.parametrize("gpus", [1, 2])
def test_batch_of_samples_equals_separate_runs(tools, gpus):
    """CONGA_BATCH: three BAMs against one CNV set / mappability track in ONE process (contexts, SONIC, BEDs and
    flattened tracks set up once, sample k+1 decoded under sample k) — each sample's files must be the bytes a
    separate run of the reference writes for it."""
    if gpus > 1:
        import torch
        if torch.cuda.device_count() < gpus:
            pytest.skip("needs %d GPUs" % gpus)
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(106)
        chroms = [("1", 300_000), ("2", 200_000), ("3", 120_000)]
        paths = cli_inputs.build_inputs(tmp, chroms, rng, with_map=True)
        bams = [paths["bam"], cli_inputs.extra_bam(tmp, chroms, rng, "b", 30000), cli_inputs.extra_bam(tmp, chroms, rng, "c", 8000)]
        flags = ["--dels", paths["dels"], "--dups", paths["dups"], "--mappability", paths["map"], "--min-mapq", "10"]
        want = []
        for i, bam in enumerate(bams):
            wd = os.path.join(tmp, "ref%d" % i)
            os.makedirs(os.path.join(wd, "o"))
            p = subprocess.run([REF_CONGA, "-i", bam, "--ref", paths["ref"], "--sonic", paths["sonic"], "--out", os.path.join(wd, "o", "s")] + flags,
                               cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
            assert p.returncode == 0
            files = {nm: open(os.path.join(wd, "o", "s_" + nm), "rb").read() for nm in ("svs.bed", "dels.bed", "dups.bed")}
            files["log"] = [l for l in open(os.path.join(wd, "conga.log"), "rb").read().split(b"\n") if l.startswith(b"Read Count")]
            want.append(files)
        wd = os.path.join(tmp, "batch")
        os.makedirs(wd)
        with open(os.path.join(wd, "batch.tsv"), "w") as f:
            for i, bam in enumerate(bams):
                f.write("%s\t%s\n" % (bam, os.path.join(wd, "s%d" % i)))
        p = subprocess.run([HOST, "-i", bams[0], "--ref", paths["ref"], "--sonic", paths["sonic"], "--out", os.path.join(wd, "unused")] + flags,
                           cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600,
                           env=dict(os.environ, CONGA_BATCH=os.path.join(wd, "batch.tsv"), CONGA_GPUS=str(gpus)))
        assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
        for i in range(3):
            for nm in ("svs.bed", "dels.bed", "dups.bed"):
                assert open(os.path.join(wd, "s%d_%s" % (i, nm)), "rb").read() == want[i][nm], "sample %d %s" % (i, nm)
            log = [l for l in open(os.path.join(wd, "s%d_conga.log" % i), "rb").read().split(b"\n") if l.startswith(b"Read Count")]
            assert log == want[i]["log"]
        assert want[0]["dels.bed"] != want[1]["dels.bed"]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.parametrize("case", ["mq20", "map_default_mq", "ragged_length", "repeats_mq20", "repeats_mq_default"])
def test_split_read_support(tools, case):
    """--rp with --dups (BASELINE.json configs[3]): the split-read side path — gate, average-phred tests, 10-mer
    index of the chromosome, re-mapping of both read halves within +-100 kb at <= 5 % mismatches, FORWARD/FORWARD
    pairing, satellite filter, +-50 widening, overlap counts — must reproduce the reference's READ_PAIR columns
    and the read-pair rule for duplications in _svs.bed, byte for byte."""
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng({"mq20": 201, "map_default_mq": 202, "ragged_length": 203, "repeats_mq20": 205, "repeats_mq_default": 206}[case])
        if case == "mq20":
            paths = cli_inputs.build_inputs_rp(tmp, [("1", 400_000), ("2", 250_000)], rng)
            flags = ["--rp", "5", "--min-mapq", "20"]
        elif case == "map_default_mq":
            paths = cli_inputs.build_inputs_rp(tmp, [("1", 300_000), ("chr2", 200_000), ("X", 60_000)], rng, with_map=True)
            flags = ["--rp", "12", "--mappability", paths["map"], "--min-read-length", "70"]
        elif case.startswith("repeats"):                         # tandem repeats (mapq = 60 / n), >= 100 hits, inverted repeats, Ns, chromosome ends
            paths = cli_inputs.build_inputs_rp(tmp, [("1", 350_000), ("2", 220_000)], rng, repeats=True)
            flags = ["--rp", "4"] + (["--min-mapq", "20"] if case == "repeats_mq20" else [])
        else:                                                    # lengths that are not multiples of 1000: the tail is never indexed
            paths = cli_inputs.build_inputs_rp(tmp, [("1", 300_700), ("2", 150_300)], rng, junction_reads=8)
            flags = ["--rp", "3", "--min-mapq", "9", "--c-score", "0.4"]
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--dups", paths["dups"]] + flags)
        assert_same(outs)
        read_pair = [int(l.split(b"\t")[5]) for l in outs["ref"]["s_dups.bed"].split(b"\n")[1:] if l]
        border = [int(l.split(b"\t")[5]) for l in outs["ref"]["s_dels.bed"].split(b"\n")[1:] if l]
        assert max(read_pair) > 0 and max(border) > 0, "no read-pair support in the case: not a test"
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def test_rp_without_dups_changes_nothing(tools):
    """--rp with only --dels leaves the split-read machinery off (bam_data.c:207: `!no_sr && dup_file`)."""
    tmp = tempfile.mkdtemp(prefix="cli_")
    try:
        rng = np.random.default_rng(204)
        paths = cli_inputs.build_inputs_rp(tmp, [("1", 200_000)], rng, junction_reads=5)
        outs = run_both(tmp, paths, ["--dels", paths["dels"], "--rp", "5"])
        assert_same(outs, expect_rows=False)
        assert outs["ref"]["s_dels.bed"].count(b"\n") > 5
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


@pytest.mark.parametrize("workload", ["c1", "c2", "c3"])
def test_baseline_configs_at_full_size_whole_program(tools, workload):
    """BASELINE.json configs[0], configs[1] and configs[2] at their full sizes (C2: 22 GRCh37-length chromosomes, 4.8 M
    reads, 10 k deletions + 5 k duplications, mappability BED; C3: 43 M reads at 1x, 33 k deletions + 6 k duplications —
    about 75 s, half of it building the BAM): every byte of the three BED files and every conga.log
    chromosome line equals what the unmodified reference program writes from the same BAM."""
    import json
    import sys
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "cli_bench.py"), "--workload", workload, "--runs", "1"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=900)
    assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
    res = json.loads(p.stdout.decode().strip().splitlines()[-1])
    assert res["identical_outputs"] is True
    assert res["rows"]["s_dels.bed"] > 100 and res["rows"]["s_svs.bed"] > 10
