// Drop-in replacements for command_union / command_sub (RabbitKSSD src/subCommand.h:11-12, implemented in
// src/subCommand.cpp:307-543, :545-794) on top of libkssd_b200.so.
//
// Same signatures and files: `union` writes ONE sketch named "<sketchFile> merged sketches" that holds every hash of every
// genome once, ascending (the order the reference's bitmap scan produces); `sub` writes the query sketches without the
// hashes that occur in any reference sketch.  The reference streams the input file through a producer / consumer queue
// into per-thread bitmaps over the whole hash space (512 MB each for 32-bit hashes); here the file is read with the
// reference's own readSketches(), the set arithmetic runs on the GPU (sort + unique; binary search + compaction) and the
// result is written with the reference's own saveSketches().  `sub` keeps the query genomes in file order (the reference
// writes them in the order its consumer threads finish) and the surviving hashes in their original order.
#include <err.h>

#include <cerrno>
#include <cstring>
#include <vector>

#include "kssd_host.h"
#include "sketch.h"
#include "subCommand.h"

namespace {

using kssd_host::check;

// all hashes of all genomes, flat (+ CSR offsets)
template <typename H>
void flatten(const std::vector<sketch_t> &sk, bool use64, std::vector<H> &flat, std::vector<uint64_t> &off) {
    off.assign(sk.size() + 1, 0);
    for (size_t i = 0; i < sk.size(); i++) off[i + 1] = off[i] + (use64 ? sk[i].hashSet64.size() : sk[i].hashSet.size());
    flat.resize(off.back());
    for (size_t i = 0; i < sk.size(); i++) {
        const size_t n = off[i + 1] - off[i];
        if (!n) continue;
        if (use64) memcpy(flat.data() + off[i], sk[i].hashSet64.data(), n * sizeof(H));
        else memcpy(flat.data() + off[i], sk[i].hashSet.data(), n * sizeof(H));
    }
}

template <typename H>
std::vector<H> union_of(kssd_ctx *ctx, const std::vector<sketch_t> &sk, bool use64) {
    std::vector<H> flat;
    std::vector<uint64_t> off;
    flatten<H>(sk, use64, flat, off);
    const void *set = nullptr;
    uint64_t n = 0;
    check(kssd_b200_set_union(ctx, flat.data(), flat.size(), (int)sizeof(H), &set, &n), "set_union");
    const H *p = static_cast<const H *>(set);
    return std::vector<H>(p, p + n);
}

}  // namespace

void command_union(string sketchFile, string outputFile, int threads) {
    (void)threads;
    if (!isSketchFile(sketchFile)) err(errno, "ERROR: command_union, %s is not sketch file, need input sketch file\n", sketchFile.c_str());
    std::vector<sketch_t> sketches;
    sketchInfo_t info;
    readSketches(sketches, info, sketchFile);
    const bool use64 = info.half_k - info.drlevel > 8;
    std::cerr << (use64 ? "-----use hash64" : "-----not use hash64") << std::endl;
    std::cerr << "the total genome number in sketch file is: " << sketches.size() << std::endl;
    kssd_ctx *ctx = kssd_host::context();
    sketch_t s;
    s.id = 0;
    s.fileName = sketchFile + " merged sketches";  // subCommand.cpp:364
    if (use64) s.hashSet64 = union_of<uint64_t>(ctx, sketches, true);
    else s.hashSet = union_of<uint32_t>(ctx, sketches, false);
    std::vector<sketch_t> merged(1, s);
    info.genomeNumber = 1;
    saveSketches(merged, info, outputFile);
}

void command_sub(string refSketchFile, string querySketchFile, string outputFile, int threads) {
    (void)threads;
    if (!isSketchFile(refSketchFile)) err(errno, "ERROR: command_sub(), %s is not sketch file, need input sketch file\n", refSketchFile.c_str());
    if (!isSketchFile(querySketchFile)) err(errno, "ERROR: command_sub(), %s is not sketch file, need input sketch file\n", querySketchFile.c_str());
    std::vector<sketch_t> refs, queries;
    sketchInfo_t info, qinfo;
    readSketches(refs, info, refSketchFile);
    readSketches(queries, qinfo, querySketchFile);
    if (qinfo.id != info.id) {
        std::cerr << "ERROR: command_sub(), the reference and query sketches were made with different parameters" << std::endl;
        exit(1);
    }
    const bool use64 = info.half_k - info.drlevel > 8;
    kssd_ctx *ctx = kssd_host::context();
    const void *out = nullptr;
    const uint64_t *ooff = nullptr;
    if (use64) {
        const std::vector<uint64_t> ref_set = union_of<uint64_t>(ctx, refs, true);
        std::vector<uint64_t> flat;
        std::vector<uint64_t> off;
        flatten<uint64_t>(queries, true, flat, off);
        check(kssd_b200_set_subtract(ctx, flat.data(), off.data(), (uint32_t)queries.size(), ref_set.data(), ref_set.size(), 8, &out, &ooff), "set_subtract");
        const uint64_t *p = static_cast<const uint64_t *>(out);
        for (size_t i = 0; i < queries.size(); i++) queries[i].hashSet64.assign(p + ooff[i], p + ooff[i + 1]);
    } else {
        const std::vector<uint32_t> ref_set = union_of<uint32_t>(ctx, refs, false);
        std::vector<uint32_t> flat;
        std::vector<uint64_t> off;
        flatten<uint32_t>(queries, false, flat, off);
        check(kssd_b200_set_subtract(ctx, flat.data(), off.data(), (uint32_t)queries.size(), ref_set.data(), ref_set.size(), 4, &out, &ooff), "set_subtract");
        const uint32_t *p = static_cast<const uint32_t *>(out);
        for (size_t i = 0; i < queries.size(); i++) queries[i].hashSet.assign(p + ooff[i], p + ooff[i + 1]);
    }
    qinfo.genomeNumber = (int)queries.size();
    saveSketches(queries, qinfo, outputFile);
}
