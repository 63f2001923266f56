// Drop-in replacements for sketchFastaFile / sketchFastqFile / transSketches (RabbitKSSD src/sketch.h:62-66,
// implemented in src/sketch.cpp:318-593, :596-890, :894-1021) on top of libkssd_b200.so.
//
// Same signatures, same side effects (fills `sketches`, sets `info`, writes <out>.sketch and, unless isQuery,
// <out>.sketch.dict / .index), same parsers (kseq over zlib, src/kseq.h) -- but the per-base scan, the per-genome
// set / count containers and the index inversion run on the GPU.  Parser threads only strip FASTA/FASTQ syntax:
// they copy the sequence (and quality) bytes of whole records, each followed by one separator byte, into
// pinned double-buffered batches (kssd_b200_batch_acquire / _submit).
//
// Differences a caller can observe (all inside what the reference leaves unspecified):
//  * sketches come out in id order (the reference pushes them in thread-completion order, sketch.cpp:558-560);
//  * hashSet is sorted ascending (the reference stores unordered_set iteration order, sketch.cpp:537-550);
//  * files are not split into a "big" RabbitFX group and a "small" kseq group (sketch.cpp:352-450): every file is
//    read with kseq; ids are still the position in the size-descending file list (sketch.cpp:365, :556).
//  * half_k - drlevel > 8 (64-bit hashes, sketch.cpp:336): hashSet64 is filled instead, sorted ascending, and the sparse
//    .index / .dict list the hashes in ascending order (the reference writes its hash map's iteration order).
#include <err.h>
#include <fcntl.h>
#include <unistd.h>
#include <omp.h>
#include <sys/stat.h>
#include <zlib.h>

#include <algorithm>
#include <cerrno>
#include <condition_variable>
#include <cstring>
#include <fstream>
#include <mutex>
#include <vector>

#include "kseq.h"
#include "kssd_host.h"
#include "raw_slicer.h"
#include "sketch.h"

// src/sketch.cpp:17 instantiates the same static kseq functions in its own translation unit
KSEQ_INIT(gzFile, gzread)

// one per program: starts creating the CUDA context of the first GPU while main() is still parsing its arguments
static kssd_host::EarlyInit g_early_init;

namespace {

using kssd_host::check;
using kssd_host::RawSlicer;

const unsigned char kSeparator = '\n';  // any non-ACGT byte resets the valid-run counter like a new record does

// Packs pieces (whole records of one genome, separator-terminated) into the library's pinned batches.  A parser thread
// RESERVES a slice of the current batch under the lock and copies into it outside the lock, so the copies of all parser
// threads run in parallel; a batch is submitted once its last writer has finished.
class BatchFiller {
  public:
    BatchFiller(kssd_ctx *ctx, bool has_qual, int raw = 0) : ctx_(ctx), has_qual_(has_qual), raw_(raw) {}  // raw: 0 parsed records, 1 FASTA text, KSSD_RAW_FASTQ

    // RAW FASTA text (kssd_batch.raw_fasta, tokenised on the device): reserves a piece of prefix_len + n bytes (padded with
    // '\n' to a multiple of 32), copies the prefix and lets `fill` put the n text bytes in place -- outside the lock.
    template <typename Fill>
    void append_raw(uint32_t genome, const unsigned char *prefix, size_t prefix_len, size_t n, int flags, Fill &&fill) {
        const size_t need = (prefix_len + n + 31) & ~(size_t)31;
        kssd_batch *b;
        size_t at;
        const double t_in = get_sec();
        {
            std::unique_lock<std::mutex> lock(mu_);
            while (cur_ && (cur_->n_bytes + need > cur_->capacity || cur_->n_seg == cur_->seg_capacity)) submit(lock);  // (another thread may have replaced it meanwhile)
            if (!cur_) {
                check(kssd_b200_batch_acquire(ctx_, &cur_, std::max(kBatchBytes, need), 1 << 16), "batch_acquire");
                cur_->n_bytes = 0;
                cur_->n_seg = 0;
                cur_->first_emit = 0;
                cur_->raw_fasta = raw_;
            }
            b = cur_;
            at = b->n_bytes;
            b->seg_off[b->n_seg] = at;
            b->seg_genome[b->n_seg] = genome;
            b->seg_flags[b->n_seg] = (uint8_t)flags;
            b->seg_first[b->n_seg] = (uint32_t)prefix_len;
            b->n_seg++;
            b->n_bytes += need;
            b->seg_off[b->n_seg] = b->n_bytes;
            writers_++;
        }
        const double t_got = get_sec();
        if (prefix_len) memcpy(b->seq + at, prefix, prefix_len);
        fill(b->seq + at + prefix_len);
        memset(b->seq + at + prefix_len + n, '\n', need - prefix_len - n);
        const double t_filled = get_sec();
        {
            std::lock_guard<std::mutex> lock(mu_);
            writers_--;
            t_reserve_ += t_got - t_in;
            t_fill_ += t_filled - t_got;
        }
        cv_.notify_all();
    }

    // RAW FASTQ text (kssd_batch.raw_fasta = KSSD_RAW_FASTQ, tokenised on the device): reserves n_max bytes rounded up to 32 and
    // lets `fill` put text there -- outside the lock; fill returns how many of the bytes are WHOLE RECORDS (ending with '\n'), the
    // rest of the slice becomes 0x00 padding.  Returns that count.
    template <typename Fill>
    size_t append_raw_records(uint32_t genome, size_t n_max, Fill &&fill) {
        const size_t need = (n_max + 31) & ~(size_t)31;
        kssd_batch *b;
        size_t at;
        const double t_in = get_sec();
        {
            std::unique_lock<std::mutex> lock(mu_);
            while (cur_ && (cur_->n_bytes + need > cur_->capacity || cur_->n_seg == cur_->seg_capacity)) submit(lock);
            if (!cur_) {
                check(kssd_b200_batch_acquire(ctx_, &cur_, std::max(kBatchBytes, need), 1 << 16), "batch_acquire");
                cur_->n_bytes = 0;
                cur_->n_seg = 0;
                cur_->first_emit = 0;
                cur_->raw_fasta = raw_;
            }
            b = cur_;
            at = b->n_bytes;
            b->seg_off[b->n_seg] = at;
            b->seg_genome[b->n_seg] = genome;
            b->seg_flags[b->n_seg] = 0;
            b->seg_first[b->n_seg] = 0;
            b->n_seg++;
            b->n_bytes += need;
            b->seg_off[b->n_seg] = b->n_bytes;
            writers_++;
        }
        const double t_got = get_sec();
        const size_t used = std::min(fill(b->seq + at), n_max);
        memset(b->seq + at + used, 0, need - used);
        const double t_filled = get_sec();
        {
            std::lock_guard<std::mutex> lock(mu_);
            writers_--;
            t_reserve_ += t_got - t_in;
            t_fill_ += t_filled - t_got;
        }
        cv_.notify_all();
        return used;
    }

    // seq[0, n) (+ qual) of `genome`; when `terminate` one separator byte is appended after the bytes
    void append(uint32_t genome, const unsigned char *seq, const unsigned char *qual, size_t n, bool terminate) {
        const size_t need = n + (terminate ? 1 : 0);
        if (need == 0) return;
        kssd_batch *b;
        size_t at;
        {
            std::unique_lock<std::mutex> lock(mu_);
            while (cur_ && (cur_->n_bytes + need > cur_->capacity || cur_->n_seg == cur_->seg_capacity)) submit(lock);  // (another thread may have replaced it meanwhile)
            if (!cur_) {
                check(kssd_b200_batch_acquire(ctx_, &cur_, std::max(kBatchBytes, need), 1 << 16), "batch_acquire");
                cur_->n_bytes = 0;
                cur_->n_seg = 0;
                cur_->first_emit = 0;
            }
            b = cur_;
            at = b->n_bytes;
            b->seg_off[b->n_seg] = at;
            b->seg_genome[b->n_seg] = genome;
            b->n_seg++;
            b->n_bytes += need;
            b->seg_off[b->n_seg] = b->n_bytes;
            writers_++;
        }
        memcpy(b->seq + at, seq, n);
        if (terminate) b->seq[at + n] = kSeparator;
        if (has_qual_) {
            if (qual) memcpy(b->qual + at, qual, n);
            else memset(b->qual + at, 0x7F, n);
            if (terminate) b->qual[at + n] = 0x7F;
        }
        {
            std::lock_guard<std::mutex> lock(mu_);
            writers_--;
        }
        cv_.notify_all();
    }

    void flush() {
        std::unique_lock<std::mutex> lock(mu_);
        if (cur_) submit(lock);
        if (getenv("KSSD_B200_TRACE") && (t_fill_ > 0 || t_submit_ > 0))
            std::cerr << "[kssd_b200 trace] filler(" << (raw_ == KSSD_RAW_FASTQ ? "raw fastq" : raw_ ? "raw" : "parsed") << "): thread-seconds waiting to reserve " << t_reserve_ << ", copying " << t_fill_
                      << "; submits " << n_submit_ << " took " << t_submit_ << " s, of which waiting for writers " << t_drain_ << std::endl;
    }

    static constexpr size_t kBatchBytes = 64u << 20;   // pinned batch size: small enough that the first copy starts early and page-locking is cheap
    static constexpr size_t kPieceBytes = 8u << 20;    // small records are gathered per thread up to this much
    static constexpr size_t kDirectBytes = 1u << 20;   // records at least this long go straight from the parser's buffer

  private:
    // The current batch is full: wait for the copies into its reserved slices, then hand it to the library.  The wait
    // releases the lock, so several threads can arrive here for the SAME batch: the first one to see it drained submits it,
    // the others notice that cur_ has moved on and return (they must not submit the next, half-empty batch).
    void submit(std::unique_lock<std::mutex> &lock) {
        const double t0 = get_sec();
        const uint64_t mine = gen_;
        cv_.wait(lock, [&] { return writers_ == 0 || gen_ != mine; });
        if (gen_ != mine) return;
        const double t1 = get_sec();
        if (cur_->n_bytes) check(kssd_b200_batch_submit(ctx_, cur_), "batch_submit");
        cur_ = nullptr;
        gen_++;
        cv_.notify_all();
        t_drain_ += t1 - t0;
        t_submit_ += get_sec() - t0;
        n_submit_++;
    }
    uint64_t gen_ = 0;  // batches submitted so far
    double t_reserve_ = 0, t_fill_ = 0, t_submit_ = 0, t_drain_ = 0;
    int n_submit_ = 0;
    kssd_ctx *ctx_;
    bool has_qual_;
    int raw_;
    std::mutex mu_;
    std::condition_variable cv_;
    int writers_ = 0;  // slices reserved in cur_ whose copy has not finished
    kssd_batch *cur_ = nullptr;
};

// ---- plain FASTA files: the text goes to the GPU as it is ----------------------------------------
// The host only reads bytes: a file is cut into pieces of kRawPiece bytes, each read straight into the pinned batch, and the
// device strips headers and line ends (fasta.cu).  What the host keeps per file is the line state at the cut (inside a header
// line / at a line start / inside a sequence line) and the last kCtx sequence bytes of the running record, which prefix the
// next piece so that k-mers across the cut are seen (kssd_batch.raw_fasta).  One thread reads one file front to back -- a
// 3 Gbp genome moves at memory-copy speed instead of through kseq's byte loop (the reference uses RabbitFX's chunked
// reader for such files, src/sketch.cpp:173-238, :380-450).
// Piece size: a batch (64 MiB) takes reservations from all parser threads at once, but only until it is full -- with 16 MiB
// pieces three or four threads filled a batch while the others waited for the next one (20 GB of FASTQ moved at 4.9 GB/s,
// profiles/r02x); 2 MiB gives every thread a slice in the open batch.
size_t raw_piece_bytes() {
    const char *e = getenv("KSSD_B200_RAW_PIECE_BYTES");  // tests cut small files into many pieces
    const size_t v = e ? (size_t)strtoull(e, nullptr, 10) : (size_t)(2u << 20);
    return std::max<size_t>(v, 64);
}

// plain FASTA = not gzip (magic 1f 8b) and the first byte is '>' (kseq skips anything before the first '>' / '@': such files
// and FASTQ go through kseq)
bool looks_like_plain_fasta(int fd) {
    unsigned char b[2] = {0, 0};
    return pread(fd, b, 2, 0) >= 1 && b[0] == '>';
}

// -> false on a read error (the caller falls back to kseq for the whole job)
bool read_raw_fasta(int fd, uint64_t size, uint32_t genome, BatchFiller &filler) {
    const size_t piece = raw_piece_bytes();
    RawSlicer sl;
    uint64_t pos = 0;
    bool ok = true;
    while (pos < size && ok) {
        uint64_t end = std::min<uint64_t>(size, pos + piece);
        if (end < size) {  // a piece must not end with the '\r' of a "\r\n"
            unsigned char last = 0;
            if (pread(fd, &last, 1, (off_t)(end - 1)) == 1 && last == '\r' && end - 1 > pos) end--;
        }
        unsigned char prefix[RawSlicer::kCtx + 1];
        size_t plen = 0;
        const int flags = pos == 0 ? 0 : sl.next_prefix(prefix, plen);
        const size_t n = (size_t)(end - pos);
        unsigned char next = '\n';
        if (end < size && pread(fd, &next, 1, (off_t)end) != 1) next = '\n';
        filler.append_raw(genome, prefix, plen, n, flags, [&](unsigned char *dst) {
            size_t got = 0;
            while (got < n) {
                const ssize_t r = pread(fd, dst + got, n - got, (off_t)(pos + got));
                if (r <= 0) { ok = false; memset(dst + got, '\n', n - got); break; }
                got += (size_t)r;
            }
            if (end < size) sl.advance(dst, n, next);
        });
        pos = end;
    }
    return ok;
}

// plain four-line FASTQ = not gzip and the first byte is '@'
bool looks_like_plain_fastq(int fd) {
    unsigned char b[2] = {0, 0};
    return pread(fd, b, 2, 0) >= 1 && b[0] == '@';
}

// Plain FASTQ files: the text goes to the GPU as it is, in pieces of whole records (fastq.cu).  Reads never share k-mers
// (src/sketch.cpp:273-283 restarts the run at every read), so the host keeps no context: it only finds the last fourth line
// end of what it read -- one pass that counts '\n' -- and starts the next piece there.
// -> false when the file cannot go this way (read error, a record longer than a piece, an incomplete last record): the caller
// falls back to kseq for the whole job.
bool read_raw_fastq(int fd, uint64_t size, uint32_t genome, BatchFiller &filler) {
    const size_t piece = std::max<size_t>(raw_piece_bytes(), 4096);
    uint64_t pos = 0;
    while (pos < size) {
        const size_t n = (size_t)std::min<uint64_t>(size - pos, piece);
        const bool last = pos + n == size;
        bool ok = true;
        const size_t used = filler.append_raw_records(genome, n + 1, [&](unsigned char *dst) -> size_t {
            size_t got = 0;
            while (got < n) {
                const ssize_t r = pread(fd, dst + got, n - got, (off_t)(pos + got));
                if (r <= 0) { ok = false; return 0; }
                got += (size_t)r;
            }
            size_t len = n;
            if (last && dst[len - 1] != '\n') dst[len++] = '\n';  // the last line may lack its line end
            const size_t whole = kssd_host::fastq_whole_records(dst, len);
            if (last)  // what follows the last whole record must be nothing kseq would read (it skips to the next '@' / '>')
                for (size_t i = whole; i < len; i++)
                    if (dst[i] == '@' || dst[i] == '>') ok = false;
            return whole;
        });
        if (!ok || used == 0) return false;
        if (last) break;
        pos += used;
    }
    return true;
}

bool by_size_desc(const fileInfo_t &a, const fileInfo_t &b) { return a.fileSize > b.fileSize; }  // cmpFile, sketch.cpp:39

bool sketch_files(const std::string &inputFile, bool isQuery, int numThreads, const kssd_parameter_t &p, bool fastq,
                  int leastQual, int leastNumKmer, std::vector<sketch_t> &sketches, sketchInfo_t &info, std::string outputFile,
                  bool allow_raw = true) {
    if (getenv("KSSD_B200_NO_RAW_FASTA")) allow_raw = false;
    const bool use64 = p.half_k - p.drlevel > 8;  // sketch.cpp:336
    // file list: as sketch.cpp:348-374 (stat every file, sort by size descending)
    std::ifstream fs(inputFile);
    if (!fs) err(errno, "cannot open the inputFile: %s\n", inputFile.c_str());
    std::vector<fileInfo_t> fileList;
    std::string fileName;
    while (getline(fs, fileName)) {
        struct stat st;
        st.st_size = 0;
        stat(fileName.c_str(), &st);
        fileInfo_t f;
        f.fileName = fileName;
        f.fileSize = st.st_size;
        fileList.push_back(f);
    }
    std::sort(fileList.begin(), fileList.end(), by_size_desc);
    const size_t n = fileList.size();
    std::cerr << "the total fileNumber is: " << n << std::endl;

    const bool trace = getenv("KSSD_B200_TRACE") != nullptr;  // phase times of the shim on stderr
    double tr0 = get_sec();
    auto lap = [&](const char *what) {
        if (!trace) return;
        const double t = get_sec();
        std::cerr << "[kssd_b200 trace] " << what << ": " << t - tr0 << " s" << std::endl;
        tr0 = t;
    };
    // one context per GPU; file t (position in the size-descending list) goes to GPU t % G, which balances the bytes
    // one GPU takes text at PCIe speed (55 GB/s), far more than parser threads deliver: another GPU only per 64 GB of input
    double total_bytes = 0;
    for (const fileInfo_t &f : fileList) total_bytes += (double)f.fileSize;
    const size_t G = std::max<size_t>(1, std::min(kssd_host::gpus_for(total_bytes, 64e9, "KSSD_B200_SKETCH_BYTES_PER_GPU"), n));
    std::vector<kssd_ctx *> ctxs = kssd_host::contexts_up_to(G);
    lap("contexts");
#pragma omp parallel for num_threads((int)G) schedule(static, 1)
    for (size_t g = 0; g < G; g++) {  // the table build walks all 16^half_subk entries: one host thread per context
        check(kssd_b200_set_params(ctxs[g], p.half_k, p.half_subk, p.drlevel, p.shuffled_dim), "set_params");
        check(kssd_b200_sketch_begin(ctxs[g], (uint32_t)n, fastq ? 1 : 0, leastQual, leastNumKmer), "sketch_begin");
    }
    std::vector<BatchFiller *> fillers(G), raw_fillers(G);
    for (size_t g = 0; g < G; g++) {
        fillers[g] = new BatchFiller(ctxs[g], fastq);
        raw_fillers[g] = new BatchFiller(ctxs[g], fastq, fastq ? KSSD_RAW_FASTQ : 1);
    }
    bool raw_failed = false;
    lap("set_params + sketch_begin");

    if (numThreads < 1) numThreads = 1;
#pragma omp parallel for num_threads(numThreads) schedule(dynamic)
    for (size_t t = 0; t < n; t++) {
        if (allow_raw) {  // plain FASTA / four-line FASTQ text: straight to the GPU, tokenised there (fasta.cu, fastq.cu)
            const int fd = open(fileList[t].fileName.c_str(), O_RDONLY);
            if (fd < 0) err(errno, "cannot open the genome file: %s\n", fileList[t].fileName.c_str());
            struct stat st;
            const bool plain = fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0 && (fastq ? looks_like_plain_fastq(fd) : looks_like_plain_fasta(fd));
            if (plain && !(fastq ? read_raw_fastq(fd, (uint64_t)st.st_size, (uint32_t)t, *raw_fillers[t % G])
                                 : read_raw_fasta(fd, (uint64_t)st.st_size, (uint32_t)t, *raw_fillers[t % G]))) {
#pragma omp atomic write
                raw_failed = true;
            }
            close(fd);
            if (plain) continue;
        }
        gzFile fp = gzopen(fileList[t].fileName.c_str(), "r");
        if (fp == NULL) err(errno, "cannot open the genome file: %s\n", fileList[t].fileName.c_str());
        gzbuffer(fp, 1 << 20);
        kseq_t *ks = kseq_init(fp);
        BatchFiller &filler = *fillers[t % G];
        std::vector<unsigned char> seq, qual;  // small records of this genome, gathered
        while (true) {
            const int length = kseq_read(ks);  // int, as in the reference (sketch.cpp:475): one record < 2^31 bases
            if (length < 0) break;
            const size_t ql = fastq && ks->qual.s ? std::min<size_t>(ks->qual.l, (size_t)length) : 0;
            if ((size_t)length >= BatchFiller::kDirectBytes && (!fastq || ql == (size_t)length)) {
                if (!seq.empty()) {
                    filler.append((uint32_t)t, seq.data(), fastq ? qual.data() : nullptr, seq.size(), false);
                    seq.clear();
                    qual.clear();
                }
                filler.append((uint32_t)t, (const unsigned char *)ks->seq.s, fastq ? (const unsigned char *)ks->qual.s : nullptr, (size_t)length, true);
                continue;
            }
            const size_t at = seq.size();
            seq.resize(at + (size_t)length + 1);
            memcpy(seq.data() + at, ks->seq.s, (size_t)length);
            seq[at + length] = kSeparator;
            if (fastq) {
                qual.resize(at + (size_t)length + 1);
                if (ql) memcpy(qual.data() + at, ks->qual.s, ql);
                memset(qual.data() + at + ql, 0x7F, (size_t)length + 1 - ql);  // no quality string: nothing is filtered
            }
            if (seq.size() >= BatchFiller::kPieceBytes) {
                filler.append((uint32_t)t, seq.data(), fastq ? qual.data() : nullptr, seq.size(), false);
                seq.clear();
                qual.clear();
            }
        }
        if (!seq.empty()) filler.append((uint32_t)t, seq.data(), fastq ? qual.data() : nullptr, seq.size(), false);
        kseq_destroy(ks);
        gzclose(fp);
    }

    lap("read / parse files into batches");
    std::vector<const uint32_t *> hashes(G, nullptr);
    std::vector<const uint64_t *> hashes64(G, nullptr), off(G, nullptr);
    std::vector<int> rcs(G, KSSD_OK);
#pragma omp parallel for num_threads((int)G) schedule(static, 1)
    for (size_t g = 0; g < G; g++) {
        fillers[g]->flush();
        raw_fillers[g]->flush();
        rcs[g] = use64 ? kssd_b200_sketch_finish64(ctxs[g], &hashes64[g], &off[g]) : kssd_b200_sketch_finish(ctxs[g], &hashes[g], &off[g]);
        if (rcs[g] != KSSD_OK && rcs[g] != KSSD_E_SYNTAX) check(rcs[g], "sketch_finish");
        delete fillers[g];
        delete raw_fillers[g];
    }
    for (size_t g = 0; g < G; g++) raw_failed |= rcs[g] == KSSD_E_SYNTAX;
    if (raw_failed) {  // text that is not plain FASTA / four-line FASTQ (or a read error): the host parser decides, as in the reference
        std::cerr << "-----raw text was not accepted by the device tokeniser, parsing all files with kseq" << std::endl;
        return sketch_files(inputFile, isQuery, numThreads, p, fastq, leastQual, leastNumKmer, sketches, info, outputFile, false);
    }
    lap("flush + sketch_finish");
    sketches.reserve(sketches.size() + n);
    for (size_t t = 0; t < n; t++) {
        const size_t g = t % G;
        sketch_t s;
        s.fileName = fileList[t].fileName;
        s.id = (int)t;
        if (use64) s.hashSet64.assign(hashes64[g] + off[g][t], hashes64[g] + off[g][t + 1]);
        else s.hashSet.assign(hashes[g] + off[g][t], hashes[g] + off[g][t + 1]);
        sketches.push_back(s);
    }

    // tail: sketch.cpp:570-590
    if (!isSketchFile(outputFile)) outputFile = outputFile + ".sketch";
    info.half_k = p.half_k;
    info.half_subk = p.half_subk;
    info.drlevel = p.drlevel;
    info.id = (p.half_k << 8) + (p.half_subk << 4) + p.drlevel;
    info.genomeNumber = sketches.size();
    lap("sketch_t vectors");
    saveSketches(sketches, info, outputFile);
    lap("saveSketches");
    std::cerr << "save the sketches into: " << outputFile << std::endl;
    if (!isQuery) {
        const double t0 = get_sec();
        transSketches(sketches, info, outputFile + ".dict", outputFile + ".index", numThreads);
        std::cerr << "===============the time of transSketches is: " << get_sec() - t0 << std::endl;
    }
    return true;
}

}  // namespace

bool sketchFastaFile(string inputFile, bool isQuery, int numThreads, kssd_parameter_t parameter, vector<sketch_t> &sketches,
                     sketchInfo_t &info, string outputFile) {
    return sketch_files(inputFile, isQuery, numThreads, parameter, false, 0, 1, sketches, info, outputFile);
}

bool sketchFastqFile(string inputFile, bool isQuery, int numThreads, kssd_parameter_t parameter, int leastQual, int leastNumKmer,
                     vector<sketch_t> &sketches, sketchInfo_t &info, string outputFile) {
    return sketch_files(inputFile, isQuery, numThreads, parameter, true, leastQual, leastNumKmer, sketches, info, outputFile);
}

// transSketches (sketch.cpp:894-1021).  32-bit path (:970-1017): .dict = posting lists in ascending hash order, ids
// ascending inside a list; .index = { size_t hashSize; uint64_t totalIndex; uint32_t count[hashSize] }.  64-bit path
// (:904-969): .index = { size_t hash_number; uint64_t hash[hash_number]; uint32_t count[hash_number] }, .dict = the lists
// in the same order -- the reference's order is its hash map's iteration order, here it is ascending hash.
void transSketches(vector<sketch_t> &sketches, sketchInfo_t &info, string dictFile, string indexFile, int numThreads) {
    (void)numThreads;
    const bool use64 = info.half_k - info.drlevel > 8;
    std::cerr << (use64 ? "transSketches: use64" : "transSketches: not use64") << std::endl;
    std::vector<uint64_t> off(sketches.size() + 1, 0);
    for (size_t i = 0; i < sketches.size(); i++) off[i + 1] = off[i] + (use64 ? sketches[i].hashSet64.size() : sketches[i].hashSet.size());
    kssd_ctx *ctx = kssd_host::context();
    kssd_host::IndexCache &cache = kssd_host::index_cache();
    cache.valid = false;
    kssd_host::pending_writes().wait(&indexFile);  // an earlier background write of the same file
    auto open_or_die = [](const std::string &path, const char *what) {
        FILE *f = fopen(path.c_str(), "w+");
        if (!f) {
            std::cerr << "ERROR: transSketches, cannot open " << what << ": " << path << std::endl;
            exit(1);
        }
        return f;
    };
    uint64_t n_uniq = 0, total = 0;
    const uint32_t *uniq_count = nullptr, *dict = nullptr;
    if (use64) {
        std::vector<uint64_t> flat(off.back());
#pragma omp parallel for schedule(dynamic, 64)
        for (size_t i = 0; i < sketches.size(); i++)
            if (!sketches[i].hashSet64.empty()) memcpy(flat.data() + off[i], sketches[i].hashSet64.data(), sketches[i].hashSet64.size() * sizeof(uint64_t));
        const uint64_t *uniq_hash = nullptr;
        check(kssd_b200_index_build64(ctx, flat.data(), off.data(), (uint32_t)sketches.size(), KSSD_PTR_HOST), "index_build64");
        check(kssd_b200_index_export64(ctx, &n_uniq, &total, &uniq_hash, &uniq_count, &dict), "index_export64");
        std::cerr << "the hash_number is: " << n_uniq << std::endl;
        std::cerr << "the total size is: " << total << std::endl;
        FILE *fp0 = open_or_die(dictFile, "dictFile");
        fwrite(dict, sizeof(uint32_t), total, fp0);
        fclose(fp0);
        FILE *fp1 = open_or_die(indexFile, "indexFile");
        const size_t hash_number = n_uniq;
        fwrite(&hash_number, sizeof(size_t), 1, fp1);
        fwrite(uniq_hash, sizeof(uint64_t), hash_number, fp1);
        fwrite(uniq_count, sizeof(uint32_t), hash_number, fp1);
        fclose(fp1);
        cache = kssd_host::IndexCache{true, true, indexFile, dictFile, sketches.size()};
        return;
    }
    const size_t hashSize = 1LLU << (4 * (info.half_k - info.drlevel));
    std::cerr << "the hashSize is: " << hashSize << std::endl;
    std::vector<uint32_t> flat(off.back());
#pragma omp parallel for schedule(dynamic, 64)
    for (size_t i = 0; i < sketches.size(); i++)
        if (!sketches[i].hashSet.empty()) memcpy(flat.data() + off[i], sketches[i].hashSet.data(), sketches[i].hashSet.size() * sizeof(uint32_t));

    check(kssd_b200_index_build(ctx, flat.data(), off.data(), (uint32_t)sketches.size(), KSSD_PTR_HOST), "index_build");
    const uint32_t *uniq_hash = nullptr;
    check(kssd_b200_index_export(ctx, &n_uniq, &total, &uniq_hash, &uniq_count, &dict), "index_export");

    FILE *fp0 = open_or_die(dictFile, "dictFile");
    fwrite(dict, sizeof(uint32_t), total, fp0);
    fclose(fp0);

    // The format wants the dense 2^(4(k-l))-entry count array (1 GiB for L3K10), almost all zeros.  It is produced by a
    // background thread: a distance step that follows in this process (alldist / dist on a genome list) works on the CSR
    // that is still resident on the device instead of reading the file back (kssd_host::index_cache).
    FILE *fp1 = open_or_die(indexFile, "indexFile");
    const uint64_t totalIndex = total;
    std::vector<uint32_t> uh(uniq_hash, uniq_hash + n_uniq), uc(uniq_count, uniq_count + n_uniq);
    for (uint64_t u = 0; u < n_uniq; u++)
        if (uh[u] >= hashSize) {
            std::cerr << "ERROR: transSketches, hash " << uh[u] << " outside the hash space " << hashSize << std::endl;
            exit(1);
        }
    kssd_host::pending_writes().add(indexFile, std::thread([fp1, hashSize, totalIndex, uh = std::move(uh), uc = std::move(uc)]() {
        fwrite(&hashSize, sizeof(size_t), 1, fp1);
        fwrite(&totalIndex, sizeof(uint64_t), 1, fp1);
        const size_t kChunk = 1u << 22;  // 16 MiB of counters at a time: no 1 GiB allocation
        std::vector<uint32_t> counts(std::min<size_t>(kChunk, hashSize));
        size_t u = 0;
        for (size_t h0 = 0; h0 < hashSize; h0 += kChunk) {
            const size_t h1 = std::min(hashSize, h0 + kChunk);
            std::fill(counts.begin(), counts.begin() + (h1 - h0), 0u);
            for (; u < uh.size() && uh[u] < h1; u++) counts[uh[u] - h0] = uc[u];  // distinct hashes ascend
            fwrite(counts.data(), sizeof(uint32_t), h1 - h0, fp1);
        }
        fclose(fp1);
    }));
    cache = kssd_host::IndexCache{true, false, indexFile, dictFile, sketches.size()};
}
