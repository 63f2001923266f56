"""GPU: FASTA tokenisation on the device (fasta.cu, kssd_batch.raw_fasta) against the oracle on the records a host parser
(kseq semantics, src/kseq.h:175-215) extracts from the same text: line widths from 1 to one-line records, CRLF, blank lines,
lowercase / N, many tiny records, '>' inside a sequence line, '@' record starts, pieces cut inside header lines and inside
sequence lines with the repeated context the host shim would add, and the '+' anomaly."""
import numpy as np
import pytest

import oracle
import util
from rabbitkssd_b200 import _lib, api, synth

pytestmark = pytest.mark.gpu


def fasta_text(records, widths, crlf=False, blank_every=0):
    """records: list of (header bytes without '>', uint8 sequence).  width 0 = the whole sequence on one line."""
    out = bytearray()
    eol = b"\r\n" if crlf else b"\n"
    for i, (h, s) in enumerate(records):
        out += b">" + h + eol
        w = widths[i % len(widths)]
        s = bytes(s)
        if w == 0:
            out += s + eol
        else:
            for n, a in enumerate(range(0, len(s), w)):
                out += s[a:a + w] + eol
                if blank_every and n % blank_every == blank_every - 1:
                    out += eol
    return bytes(out)


def parse_like_kseq(text):
    """Records as kseq_read returns them for plain FASTA: a line starting with '>' or '@' begins a record; the other lines,
    without their '\\n' (and without a '\\r' in front of it), are its sequence."""
    recs, cur = [], None
    for line in text.split(b"\n"):
        if line[:1] in (b">", b"@"):
            cur = bytearray()
            recs.append(cur)
            continue
        if cur is None:
            continue
        if line.endswith(b"\r"):
            line = line[:-1]
        cur += line
    return [np.frombuffer(bytes(r), dtype=np.uint8) for r in recs]


def slice_pieces(genome, text, piece, min_ctx=31):
    """Cut `text` into pieces of `piece` bytes the way the host shim does (sketch_gpu.cpp RawSlicer): one forward pass keeps
    the line state and the last min_ctx kept sequence bytes of the current record; a piece that begins inside a header line
    gets flag 1, one that continues a record is prefixed by those bytes (+ '\n' when it begins at a line start) with flag 2."""
    out, b0 = [], 0
    hdr, at_ls, ctx = False, True, bytearray()
    i = 0
    n = len(text)
    while b0 < n:
        b1 = min(n, b0 + piece)
        if b1 < n and text[b1 - 1:b1] == b"\r":
            b1 -= 1 if b1 - 1 > b0 else -1
        if b0 == 0:
            out.append((genome, text[0:b1], 0, 0))
        elif hdr and not at_ls:
            out.append((genome, text[b0:b1], 1, 0))
        elif at_ls:
            pre = bytes(ctx) + b"\n" if ctx else b""
            out.append((genome, pre + text[b0:b1], 2 if ctx else 0, len(pre)))
        else:
            out.append((genome, bytes(ctx) + text[b0:b1], 2, len(ctx)))
        while i < b1:  # forward state: exactly the tokeniser's rule for kept bytes
            c = text[i:i + 1]
            if at_ls:
                hdr = c in (b">", b"@")
                at_ls = False
                if hdr:
                    ctx = bytearray()
            if c == b"\n":
                at_ls = True
            elif not hdr and not (c == b"\r" and text[i + 1:i + 2] == b"\n"):
                ctx += c
                del ctx[:-min_ctx]
            i += 1
        b0 = b1
    return out


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def _expect(k, s, l, texts):
    genomes = [parse_like_kseq(t) for t in texts]
    return util.oracle_sketch(k, s, l, genomes)


def _run(ctx, k, s, l, n_genomes, batches):
    ctx.set_params(k, s, l, util.shuffled_dim(k, s, l))
    ctx.sketch_begin(n_genomes)
    for pieces in batches:
        ctx.submit_raw_fasta(pieces)
    h, off = ctx.sketch_finish()
    return [h[int(off[g]):int(off[g + 1])].astype(np.uint64) for g in range(n_genomes)]


def _texts():
    fam = synth.family(515, 5, 300000, 0.03, n_contigs=3)
    texts = []
    for gi, g in enumerate(fam):
        recs = [(b"g%d_c%d some comment > with @ signs" % (gi, ci), synth.sprinkle(c, 3 * gi + ci, n_runs=3, lowercase_frac=0.02)) for ci, c in enumerate(g)]
        if gi == 1:
            recs += [(b"tiny%d" % j, synth.random_bases(900 + j, 5 + 7 * j)) for j in range(40)]  # many records shorter / barely longer than a k-mer
        if gi == 2:
            recs.insert(1, (b"", np.zeros(0, np.uint8)))   # empty header + empty record
        texts.append(fasta_text(recs, widths=[[80], [61, 0, 7], [1], [0], [70]][gi], crlf=gi == 4, blank_every=3 if gi == 0 else 0))
    # a '>' and an '@' in the middle of a sequence line stay sequence; a line starting with '@' begins a record (kseq)
    t = texts[3]
    texts[3] = t[:5000] + b">" + t[5001:9000] + b"@" + t[9001:]
    return texts


def test_whole_files_as_pieces(ctx):
    texts = _texts()
    want = _expect(10, 6, 3, texts)
    got = _run(ctx, 10, 6, 3, len(texts), [[(g, t, 0, 0) for g, t in enumerate(texts)]])
    for g in range(len(texts)):
        assert len(want[g]) > 30 and np.array_equal(got[g], want[g]), g
    # one file per batch, in reverse order
    got = _run(ctx, 10, 6, 3, len(texts), [[(g, texts[g], 0, 0)] for g in reversed(range(len(texts)))])
    for g in range(len(texts)):
        assert np.array_equal(got[g], want[g]), g


@pytest.mark.parametrize("piece", [4099, 65536 + 17, 250000])
def test_sliced_files(ctx, piece):
    """Pieces of one file spread over batches and interleaved with other files; boundaries fall inside header lines, on line
    ends and inside sequence lines (every offset modulo the line width occurs)."""
    texts = _texts()
    want = _expect(10, 6, 3, texts)
    pieces = []
    for g, t in enumerate(texts):
        pieces += slice_pieces(g, t, piece + 13 * g)
    # long header lines and blank-line runs, so that cuts also fall inside a header line and right at a line start after one
    extra = b"".join(b">" + b"h" * (3000 + 700 * j) + b"\n" + bytes(synth.random_bases(70 + j, 2500)) + b"\n\n" for j in range(12))
    texts.append(extra)
    want = _expect(10, 6, 3, texts)
    pieces += slice_pieces(len(texts) - 1, extra, min(piece, 4099))
    assert any(p[2] == 2 and p[3] for p in pieces)
    assert any(p[2] == 1 for p in pieces)
    order = np.random.default_rng(piece).permutation(len(pieces))  # pieces are independent of each other
    batches = [[pieces[i] for i in order[a:a + 50]] for a in range(0, len(order), 50)]
    got = _run(ctx, 10, 6, 3, len(texts), batches)
    for g in range(len(texts)):
        assert np.array_equal(got[g], want[g]), (piece, g)


def test_other_parameters_and_mixed_batches(ctx):
    """k = 8 (G = 1 filter) and s = 7 (G = 4); raw text and parsed records of different genomes in one session."""
    texts = _texts()[:3]
    for (k, s, l) in [(8, 5, 2), (10, 7, 4)]:
        want = _expect(k, s, l, texts)
        got = _run(ctx, k, s, l, 3, [[(g, t, 0, 0) for g, t in enumerate(texts)]])
        for g in range(3):
            assert np.array_equal(got[g], want[g]), (k, g)
    k, s, l = 10, 6, 3
    want = _expect(k, s, l, texts)
    ctx.set_params(k, s, l, util.shuffled_dim(k, s, l))
    ctx.sketch_begin(3)
    ctx.submit_raw_fasta([(0, texts[0], 0, 0)])
    seq, _, off = synth.pack_batch([parse_like_kseq(texts[1])])
    ctx.submit_host(seq, off, np.asarray([1], np.uint32))
    ctx.submit_raw_fasta(slice_pieces(2, texts[2], 70001))
    h, goff = ctx.sketch_finish()
    for g in range(3):
        assert np.array_equal(h[int(goff[g]):int(goff[g + 1])].astype(np.uint64), want[g]), g


def test_plus_line_is_reported(ctx):
    """A line starting with '+' after a sequence line: kseq switches to FASTQ parsing there (src/kseq.h:201); the device
    tokeniser does not guess -- sketch_finish returns KSSD_E_SYNTAX and the host falls back to its parser."""
    t = _texts()[0]
    cut = t.find(b"\n", 3000) + 1
    bad = t[:cut] + b"+\n" + t[cut:]
    ctx.set_params(10, 6, 3, util.shuffled_dim(10, 6, 3))
    ctx.sketch_begin(1)
    ctx.submit_raw_fasta([(0, bad, 0, 0)])
    with pytest.raises(_lib.KssdError) as e:
        ctx.sketch_finish()
    assert e.value.code == -6
    # the context is usable afterwards
    got = _run(ctx, 10, 6, 3, 1, [[(0, t, 0, 0)]])
    assert np.array_equal(got[0], _expect(10, 6, 3, [t])[0])
