"""GPU: FASTQ tokenisation on the device (fastq.cu, kssd_batch.raw_fasta = KSSD_RAW_FASTQ) against the oracle on the reads and
qualities a host parser (kseq semantics, src/kseq.h:175-215) extracts from the same text: fixed and ragged read lengths (empty
reads, reads shorter than a k-mer, 31 / 32 / 33 bytes), N runs and lowercase, CRLF, quality lines that begin with '@', '+' and
'>', several records inside one 32-byte chunk, files cut into pieces at record boundaries and shuffled over batches, the -Q
and -n filters -- and the texts the tokeniser must refuse (KSSD_E_SYNTAX: the host shim then parses with kseq)."""
import numpy as np
import pytest

import util
from rabbitkssd_b200 import _lib, api, synth

pytestmark = pytest.mark.gpu

K, S, L = 10, 6, 3


def fastq_text(reads, quals, crlf=False, names=None):
    eol = b"\r\n" if crlf else b"\n"
    out = bytearray()
    for i, (r, q) in enumerate(zip(reads, quals)):
        out += b"@" + (names[i] if names else b"read%d some comment" % i) + eol + bytes(r) + eol + b"+" + (b"" if i % 3 else b"read%d" % i) + eol + bytes(q) + eol
    return bytes(out)


def make_reads(seed, lengths, n_runs=0, lowercase=0.0):
    """Reads of i.i.d. bases; quality 'I' with 3 % of the positions anywhere in '!' .. 'I', and the first quality byte of every
    fifth read one of '@', '+', '>' (the characters kseq dispatches on at a line start)."""
    reads, quals = [], []
    for i, n in enumerate(lengths):
        r = synth.random_bases(seed * 100003 + i, n)
        if n > 20 and (n_runs or lowercase):
            r = synth.sprinkle(r, seed + i, n_runs=n_runs, run_len=3, lowercase_frac=lowercase)
        x = synth.splitmix64(seed * 7919 + i, n)
        q = np.where(x % np.uint64(100) < np.uint64(3), (x >> np.uint64(8)) % np.uint64(41) + np.uint64(33), np.uint64(73)).astype(np.uint8)
        if n and i % 5 == 0:
            q[0] = b"@+>"[(i // 5) % 3]
        reads.append(r)
        quals.append(q)
    return reads, quals


def _files():
    files = []
    r, q = make_reads(1, [150] * 8000)
    files.append((r + r[:4000], q + q[:4000]))  # half of the reads twice: the -n 2 filter keeps something
    ragged = [int(x) for x in synth.splitmix64(77, 3000) % np.uint64(400)]
    ragged[5:12] = [0, 19, 20, 21, 31, 32, 33]
    files.append(make_reads(2, ragged, n_runs=1, lowercase=0.05))
    files.append(make_reads(3, [int(x) for x in 5 + synth.splitmix64(78, 20000) % np.uint64(36)]))  # several records per chunk
    texts = [fastq_text(*files[0]), fastq_text(*files[1], crlf=True), fastq_text(*files[2], names=[b""] * 20000)]
    # quality lines that start with the characters kseq dispatches on
    for t, (reads, quals) in zip(texts, files):
        assert any(len(q) and q[0] == ord("@") for q in quals) and any(len(q) and q[0] == ord("+") for q in quals)
    return files, texts


def cut_records(genome, text, records_per_piece):
    """Pieces of whole records (the host shim cuts at every fourth line end)."""
    ends, pos, n = [], 0, 0
    while True:
        pos = text.find(b"\n", pos)
        if pos < 0:
            break
        pos += 1
        n += 1
        if n % (4 * records_per_piece) == 0:
            ends.append(pos)
    if not ends or ends[-1] != len(text):
        ends.append(len(text))
    return [(genome, text[a:b]) for a, b in zip([0] + ends[:-1], ends)]


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def _run(ctx, n_genomes, batches, Q=0, n=1, k=K, s=S, l=L):
    ctx.set_params(k, s, l, util.shuffled_dim(k, s, l))
    ctx.sketch_begin(n_genomes, True, Q, n)
    for pieces in batches:
        ctx.submit_raw_fastq(pieces)
    h, off = ctx.sketch_finish()
    return [h[int(off[g]):int(off[g + 1])].astype(np.uint64) for g in range(n_genomes)]


def _expect(files, Q=0, n=1, k=K, s=S, l=L):
    return util.oracle_sketch(k, s, l, [f[0] for f in files], [f[1] for f in files], Q, n)


@pytest.mark.parametrize("Q,n", [(0, 1), (50, 1), (50, 2)])
def test_whole_files_as_pieces(ctx, Q, n):
    files, texts = _files()
    want = _expect(files, Q, n)
    got = _run(ctx, len(texts), [[(g, t) for g, t in enumerate(texts)]], Q, n)
    for g in range(len(texts)):
        assert np.array_equal(got[g], want[g]), (Q, n, g)
    assert all(len(w) > 20 for w in want[:2 - (n == 2)])


@pytest.mark.parametrize("per_piece", [1, 7, 500])
def test_pieces_shuffled_over_batches(ctx, per_piece):
    files, texts = _files()
    if per_piece == 1:
        files, texts = [(f[0][:600], f[1][:600]) for f in files], None
        texts = [fastq_text(*files[0]), fastq_text(*files[1], crlf=True), fastq_text(*files[2], names=[b""] * 600)]
    want = _expect(files, 45, 1)
    pieces = []
    for g, t in enumerate(texts):
        pieces += cut_records(g, t, per_piece + g)
    order = np.random.default_rng(per_piece).permutation(len(pieces))
    batches = [[pieces[i] for i in order[a:a + 400]] for a in range(0, len(order), 400)]
    assert len(batches) > 1 or per_piece == 500
    got = _run(ctx, len(texts), batches, 45, 1)
    for g in range(len(texts)):
        assert np.array_equal(got[g], want[g]), (per_piece, g)


def test_other_parameters_and_parsed_batches_in_one_session(ctx):
    files, texts = _files()
    for (k, s, l) in [(8, 5, 2), (10, 7, 4)]:
        want = _expect(files[:2], 38, 1, k, s, l)
        got = _run(ctx, 2, [[(0, texts[0])], [(1, texts[1])]], 38, 1, k, s, l)
        for g in range(2):
            assert np.array_equal(got[g], want[g]), (k, g)
    want = _expect(files[:2], 38, 1)
    ctx.set_params(K, S, L, util.shuffled_dim(K, S, L))
    ctx.sketch_begin(2, True, 38, 1)
    ctx.submit_raw_fastq([(0, texts[0])])
    seq, qual, off = synth.pack_batch([files[1][0]], [files[1][1]])
    ctx.submit_host(seq, off, np.asarray([1], np.uint32), qual=qual)
    h, goff = ctx.sketch_finish()
    for g in range(2):
        assert np.array_equal(h[int(goff[g]):int(goff[g + 1])].astype(np.uint64), want[g]), g


def _bad_texts():
    reads, quals = make_reads(9, [120] * 50)
    good = fastq_text(reads, quals)
    lines = good.split(b"\n")
    two_line_seq = b"\n".join(lines[:5] + [lines[5][:60], lines[5][60:]] + lines[6:])           # a multi-line record
    short_qual = b"\n".join(lines[:7] + [lines[7][:-3]] + lines[8:])                           # fewer quality bytes than bases
    half_record = good + b"@tail\nACGT\n"                                                      # not a whole number of records
    fasta_inside = b"\n".join(lines[:8] + [b">fa", lines[9], lines[10], lines[11]] + lines[12:])  # a '>' record
    no_plus = b"\n".join(lines[:10] + [b"-"] + lines[11:])
    nul = good[:2000] + b"\0" + good[2001:]
    return good, {"two_line_seq": two_line_seq, "short_qual": short_qual, "half_record": half_record, "fasta_inside": fasta_inside,
                  "no_plus": no_plus, "nul": nul}


@pytest.mark.parametrize("name", ["two_line_seq", "short_qual", "half_record", "fasta_inside", "no_plus", "nul"])
def test_refused_texts(ctx, name):
    """The tokeniser does not guess: sketch_finish returns KSSD_E_SYNTAX and the context stays usable."""
    good, bad = _bad_texts()
    ctx.set_params(K, S, L, util.shuffled_dim(K, S, L))
    ctx.sketch_begin(2, True, 0, 1)
    ctx.submit_raw_fastq([(0, good), (1, bad[name]), (0, good)])
    with pytest.raises(_lib.KssdError) as e:
        ctx.sketch_finish()
    assert e.value.code == -6
    reads, quals = make_reads(9, [120] * 50)
    got = _run(ctx, 1, [[(0, good)]])
    assert np.array_equal(got[0], _expect([(reads, quals)])[0])


def test_raw_fastq_needs_a_quality_session(ctx):
    good, _ = _bad_texts()
    ctx.set_params(K, S, L, util.shuffled_dim(K, S, L))
    ctx.sketch_begin(1)
    with pytest.raises(_lib.KssdError, match="without quality"):
        ctx.submit_raw_fastq([(0, good)])
    ctx.sketch_finish()
