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
#include <omp.h>
#include <sys/stat.h>
#include <zlib.h>

#include <algorithm>
#include <cerrno>
#include <cstring>
#include <fstream>
#include <mutex>
#include <vector>

#include "kseq.h"
#include "kssd_host.h"
#include "sketch.h"

// src/sketch.cpp:17 instantiates the same static kseq functions in its own translation unit
KSEQ_INIT(gzFile, gzread)

namespace {

using kssd_host::check;

const unsigned char kSeparator = '\n';  // any non-ACGT byte resets the valid-run counter like a new record does

// Packs pieces (whole records of one genome, separator-terminated) into the library's pinned batches.
class BatchFiller {
  public:
    BatchFiller(kssd_ctx *ctx, bool has_qual) : ctx_(ctx), has_qual_(has_qual) {}

    void append(uint32_t genome, const std::vector<unsigned char> &seq, const std::vector<unsigned char> &qual) {
        const size_t n = seq.size();
        if (n == 0) return;
        std::lock_guard<std::mutex> lock(mu_);
        if (cur_ && (cur_->n_bytes + n > cur_->capacity || cur_->n_seg == cur_->seg_capacity)) submit();
        if (!cur_) {
            check(kssd_b200_batch_acquire(ctx_, &cur_, std::max(kBatchBytes, n), 1 << 16), "batch_acquire");
            cur_->n_bytes = 0;
            cur_->n_seg = 0;
            cur_->first_emit = 0;
        }
        memcpy(cur_->seq + cur_->n_bytes, seq.data(), n);
        if (has_qual_) memcpy(cur_->qual + cur_->n_bytes, qual.data(), n);
        cur_->seg_off[cur_->n_seg] = cur_->n_bytes;
        cur_->seg_genome[cur_->n_seg] = genome;
        cur_->n_seg++;
        cur_->n_bytes += n;
        cur_->seg_off[cur_->n_seg] = cur_->n_bytes;
    }

    void flush() {
        std::lock_guard<std::mutex> lock(mu_);
        if (cur_) submit();
    }

    static constexpr size_t kBatchBytes = 256u << 20;  // pinned batch size
    static constexpr size_t kPieceBytes = 32u << 20;   // a parser thread hands over at most this much at a time

  private:
    void submit() {
        if (cur_->n_bytes) check(kssd_b200_batch_submit(ctx_, cur_), "batch_submit");
        cur_ = nullptr;
    }
    kssd_ctx *ctx_;
    bool has_qual_;
    std::mutex mu_;
    kssd_batch *cur_ = nullptr;
};

bool by_size_desc(const fileInfo_t &a, const fileInfo_t &b) { return a.fileSize > b.fileSize; }  // cmpFile, sketch.cpp:39

bool sketch_files(const std::string &inputFile, bool isQuery, int numThreads, const kssd_parameter_t &p, bool fastq,
                  int leastQual, int leastNumKmer, std::vector<sketch_t> &sketches, sketchInfo_t &info, std::string outputFile) {
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

    kssd_ctx *ctx = kssd_host::context();
    check(kssd_b200_set_params(ctx, p.half_k, p.half_subk, p.drlevel, p.shuffled_dim), "set_params");
    check(kssd_b200_sketch_begin(ctx, (uint32_t)n, fastq ? 1 : 0, leastQual, leastNumKmer), "sketch_begin");
    BatchFiller filler(ctx, fastq);

    if (numThreads < 1) numThreads = 1;
#pragma omp parallel for num_threads(numThreads) schedule(dynamic)
    for (size_t t = 0; t < n; t++) {
        gzFile fp = gzopen(fileList[t].fileName.c_str(), "r");
        if (fp == NULL) err(errno, "cannot open the genome file: %s\n", fileList[t].fileName.c_str());
        kseq_t *ks = kseq_init(fp);
        std::vector<unsigned char> seq, qual;
        while (true) {
            const int length = kseq_read(ks);  // int, as in the reference (sketch.cpp:475): one record < 2^31 bases
            if (length < 0) break;
            const size_t at = seq.size();
            seq.resize(at + (size_t)length + 1);
            memcpy(seq.data() + at, ks->seq.s, (size_t)length);
            seq[at + length] = kSeparator;
            if (fastq) {
                qual.resize(at + (size_t)length + 1);
                const size_t ql = ks->qual.s ? std::min<size_t>(ks->qual.l, (size_t)length) : 0;
                if (ql) memcpy(qual.data() + at, ks->qual.s, ql);
                memset(qual.data() + at + ql, 0x7F, (size_t)length + 1 - ql);  // no quality string: nothing is filtered
            }
            if (seq.size() >= BatchFiller::kPieceBytes) {
                filler.append((uint32_t)t, seq, qual);
                seq.clear();
                qual.clear();
            }
        }
        filler.append((uint32_t)t, seq, qual);
        kseq_destroy(ks);
        gzclose(fp);
    }
    filler.flush();

    const uint32_t *hashes = nullptr;
    const uint64_t *hashes64 = nullptr, *off = nullptr;
    if (use64) check(kssd_b200_sketch_finish64(ctx, &hashes64, &off), "sketch_finish64");
    else check(kssd_b200_sketch_finish(ctx, &hashes, &off), "sketch_finish");
    for (size_t t = 0; t < n; t++) {
        sketch_t s;
        s.fileName = fileList[t].fileName;
        s.id = (int)t;
        if (use64) s.hashSet64.assign(hashes64 + off[t], hashes64 + off[t + 1]);
        else s.hashSet.assign(hashes + off[t], hashes + off[t + 1]);
        sketches.push_back(s);
    }

    // tail: sketch.cpp:570-590
    if (!isSketchFile(outputFile)) outputFile = outputFile + ".sketch";
    info.half_k = p.half_k;
    info.half_subk = p.half_subk;
    info.drlevel = p.drlevel;
    info.id = (p.half_k << 8) + (p.half_subk << 4) + p.drlevel;
    info.genomeNumber = sketches.size();
    saveSketches(sketches, info, outputFile);
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

    uint32_t *counts = (uint32_t *)calloc(hashSize, sizeof(uint32_t));  // the format wants the dense array
    for (uint64_t u = 0; u < n_uniq; u++) {
        if (uniq_hash[u] >= hashSize) {
            std::cerr << "ERROR: transSketches, hash " << uniq_hash[u] << " outside the hash space " << hashSize << std::endl;
            exit(1);
        }
        counts[uniq_hash[u]] = uniq_count[u];
    }
    FILE *fp1 = open_or_die(indexFile, "indexFile");
    const uint64_t totalIndex = total;
    fwrite(&hashSize, sizeof(size_t), 1, fp1);
    fwrite(&totalIndex, sizeof(uint64_t), 1, fp1);
    fwrite(counts, sizeof(uint32_t), hashSize, fp1);
    fclose(fp1);
    free(counts);
}
