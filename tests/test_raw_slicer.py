"""CPU: the host half of the raw-FASTA path.  RawSlicer (csrc/host/raw_slicer.h) must cut a file into exactly the pieces
(flags, context prefix) that a plain forward state machine over the whole text gives -- it only looks at the tail of every
piece (memrchr), so one-line records, header lines at the cut, CRLF and tiny pieces are the cases that matter."""
import os
import subprocess

import pytest

import test_gpu_fasta as tf
from rabbitkssd_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("slicer") / "slicer_harness")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++14", "-I", os.path.join(ROOT, "include"),
                           "-I", os.path.join(ROOT, "rabbitkssd_b200", "csrc", "host"),
                           os.path.join(ROOT, "tests", "native", "slicer_harness.cpp"), "-o", exe])
    return exe


def test_slicer_equals_forward_state_machine(harness, tmp_path):
    texts = tf._texts()
    texts.append(b">h1\n" + b"ACGT" * 30000 + b"\nAC\nGT\n>h2 x\nA\n\n\nC\n>h3\n" + b"TTGA" * 5000 + b"\r\nGG\r\n")
    texts.append(b">only header, no newline at the end")
    texts.append(b">a\n" + bytes(synth.random_bases(1, 700)) + b"\n>" + b"x" * 5000 + b"\n" + bytes(synth.random_bases(2, 900)))
    p = str(tmp_path / "t.fa")
    for g, tx in enumerate(texts):
        with open(p, "wb") as f:
            f.write(tx)
        for piece in (64, 97, 1000, 4099, 65553, 120001):
            out = subprocess.run([harness, p, str(piece)], stdout=subprocess.PIPE, check=True).stdout.decode().splitlines()
            want = tf.slice_pieces(g, tx, piece)
            assert len(out) == len(want), (g, piece)
            for line, (_, data, flags, first) in zip(out, want):
                f_ = line.split(" ")
                p0, p1, fl, plen = int(f_[0]), int(f_[1]), int(f_[2]), int(f_[3])
                prefix = bytes.fromhex(f_[4]) if len(f_) > 4 else b""
                assert (fl, plen, prefix, p1 - p0) == (flags, first, data[:first], len(data) - first), (g, piece, p0)


def test_fastq_cuts_are_record_boundaries(harness, tmp_path):
    """fastq_whole_records: every piece of the raw FASTQ reader ends just behind a line end whose number is a multiple of four;
    the pieces tile the file up to its last whole record; a window without four line ends yields nothing."""
    import test_gpu_fastq as tq
    _, texts = tq._files()
    texts = [t[:200000] for t in texts]
    texts[0] = texts[0][:texts[0].rfind(b"\n@") + 1]  # whole records
    texts.append(texts[0] + b"\n\n")                  # two blank lines at the end
    texts.append(texts[0] + b"@x\nAC")                # an unfinished record at the end
    p = str(tmp_path / "t.fq")
    for g, tx in enumerate(texts):
        with open(p, "wb") as f:
            f.write(tx)
        ends4 = [i + 1 for n, i in enumerate([i for i, c in enumerate(tx) if c == 10], 1) if n % 4 == 0]
        for piece in (4096, 4099, 70001, 1 << 20):
            out = [tuple(int(x) for x in l.split()) for l in
                   subprocess.run([harness, p, str(piece), "fastq"], stdout=subprocess.PIPE, check=True).stdout.decode().splitlines()]
            pos = 0
            for b, e in out:
                assert b == pos and e in ends4 and e == max(x for x in ends4 if x <= min(len(tx), b + piece)), (g, piece, b)
                pos = e
            assert pos == ends4[-1], (g, piece)
    with open(p, "wb") as f:
        f.write(b"@r\nACGT\n+\nIII")  # three line ends: no whole record without the reader's final '\n'
    assert subprocess.run([harness, p, "4096", "fastq"], stdout=subprocess.PIPE, check=True).stdout.decode().split() == ["0", "0"]
