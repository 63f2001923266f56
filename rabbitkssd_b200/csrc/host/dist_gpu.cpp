// Drop-in replacements for index_tridist / index_dist (RabbitKSSD src/dist.h:35-37, implemented in
// src/dist.cpp:22-343, :429-776) on top of libkssd_b200.so.
//
// Same signatures and file side effects: <refSketchOut>.index/.dict are re-read from disk exactly as the reference
// does (dist.cpp:83-130, :488-523), results go to per-"thread" sub-files <out>.dir/<out>.<t> (+ <out>.index.<t>)
// which are concatenated under the header into <out> when they total <= 4 GiB and otherwise kept together with
// <out>.index (dist.cpp:276-336, :710-770).  The posting accumulation and the O(rows x cols) epilogue run on the GPU;
// the device returns (row, col, common) for every cell that passes a conservative threshold test and this file
// recomputes Jaccard / containment and the distance with the reference's own double expressions and libm log()
// (dist.cpp:216-232, :237-251, :608-624, :647-661), so the kept pair set and the printed %f values are identical.
#include <omp.h>
#include <sys/stat.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <queue>
#include <vector>

#include "dist.h"
#include "kssd_host.h"

namespace {

using kssd_host::check;

// .index / .dict as written by transSketches: 32-bit path (sketch.cpp:1006-1011) { hashSize, totalIndex, count[hashSize] };
// 64-bit path (sketch.cpp:951-958) { hash_number, hash[hash_number], count[hash_number] }, lists in the same order
struct IndexFiles {
    size_t hashSize = 0;
    uint64_t totalIndex = 0;
    std::vector<uint64_t> hash64;
    std::vector<uint32_t> counts, dict;
};

void read_index(const std::string &indexFile, const std::string &dictFile, const char *who, bool use64, IndexFiles &ix) {
    FILE *fp = fopen(indexFile.c_str(), "rb");
    if (!fp) {
        std::cerr << "ERROR: " << who << ", cannot open the index sketch file: " << indexFile << std::endl;
        exit(1);
    }
    if (use64) {  // dist.cpp:38-54
        size_t hash_number = 0;
        const size_t r0 = fread(&hash_number, sizeof(size_t), 1, fp);
        ix.hash64.resize(hash_number);
        ix.counts.resize(hash_number);
        const size_t r1 = fread(ix.hash64.data(), sizeof(uint64_t), hash_number, fp);
        const size_t r2 = fread(ix.counts.data(), sizeof(uint32_t), hash_number, fp);
        fclose(fp);
        if (r0 != 1 || r1 != hash_number || r2 != hash_number) {
            std::cerr << "ERROR: " << who << ", error read hash_number, hash_arr, and hash_size_arr" << std::endl;
            exit(1);
        }
        ix.hashSize = hash_number;
        ix.totalIndex = 0;
        for (uint32_t n : ix.counts) ix.totalIndex += n;
        ix.dict.resize(ix.totalIndex);
        FILE *fd = fopen(dictFile.c_str(), "rb");
        if (!fd) {
            std::cerr << "ERROR: " << who << ", cannot open dict file: " << dictFile << std::endl;
            exit(1);
        }
        const size_t r3 = fread(ix.dict.data(), sizeof(uint32_t), ix.totalIndex, fd);
        fclose(fd);
        if (r3 != ix.totalIndex) {
            std::cerr << "ERROR: " << who << ", the read hash number is not equal to the saved hash number information" << std::endl;
            exit(1);
        }
        return;
    }
    const size_t r0 = fread(&ix.hashSize, sizeof(size_t), 1, fp);
    const size_t r1 = fread(&ix.totalIndex, sizeof(uint64_t), 1, fp);
    if (r0 != 1 || r1 != 1) {
        std::cerr << "ERROR: " << who << ", error read hash_size, total_index" << std::endl;
        exit(1);
    }
    ix.counts.resize(ix.hashSize);
    const size_t r2 = fread(ix.counts.data(), sizeof(uint32_t), ix.hashSize, fp);
    fclose(fp);
    ix.dict.resize(ix.totalIndex);
    FILE *fd = fopen(dictFile.c_str(), "rb");
    if (!fd) {
        std::cerr << "ERROR: " << who << ", cannot open the dictionary sketch file: " << dictFile << std::endl;
        exit(1);
    }
    const size_t r3 = fread(ix.dict.data(), sizeof(uint32_t), ix.totalIndex, fd);
    fclose(fd);
    if (r2 != ix.hashSize || r3 != ix.totalIndex) {
        std::cerr << "ERROR: " << who << ", error read hash_size, total_index, sketch_size_arr, index_arr" << std::endl;
        exit(1);
    }
}

// sketches -> CSR; the hashes are hashSet (u32) or, in the 64-bit configuration, hashSet64 (sketch.h:21-22)
struct Flat {
    std::vector<uint32_t> h32, sizes;
    std::vector<uint64_t> h64, off;
};

void flatten(const std::vector<sketch_t> &sk, bool use64, Flat &f) {
    f.off.assign(sk.size() + 1, 0);
    f.sizes.resize(sk.size());
    for (size_t i = 0; i < sk.size(); i++) {
        f.sizes[i] = (uint32_t)(use64 ? sk[i].hashSet64.size() : sk[i].hashSet.size());
        f.off[i + 1] = f.off[i] + f.sizes[i];
    }
    if (use64) f.h64.resize(f.off.back());
    else f.h32.resize(f.off.back());
    for (size_t i = 0; i < sk.size(); i++) {
        if (!f.sizes[i]) continue;
        if (use64) memcpy(f.h64.data() + f.off[i], sk[i].hashSet64.data(), (size_t)f.sizes[i] * sizeof(uint64_t));
        else memcpy(f.h32.data() + f.off[i], sk[i].hashSet.data(), (size_t)f.sizes[i] * sizeof(uint32_t));
    }
}

int collect_hits(void *user, const kssd_hit *hits, size_t n) {
    std::vector<kssd_hit> *v = reinterpret_cast<std::vector<kssd_hit> *>(user);
    v->insert(v->end(), hits, hits + n);
    return 0;
}

// the reference's expressions, verbatim in meaning: dist.cpp:216-231 / :237-250
inline void exact_distance(int common, int size0, int size1, int kmer_size, int isContainment, double &jorc, double &dist) {
    if (!isContainment) {
        const int denom = size0 + size1 - common;
        jorc = (size0 == 0 || size1 == 0) ? 0.0 : (double)common / denom;
        if (jorc == 1.0) dist = 0.0;
        else if (jorc == 0.0) dist = 1.0;
        else dist = (double)-1.0 / kmer_size * log((2 * jorc) / (1.0 + jorc));
    } else {
        const int denom = std::min(size0, size1);
        jorc = (size0 == 0 || size1 == 0) ? 0.0 : (double)common / denom;
        if (jorc == 1.0) dist = 0.0;
        else if (jorc == 0.0) dist = 1.0;
        else dist = (double)-1.0 / kmer_size * log(jorc);
    }
}

// Sub-file policy shared by both entry points (dist.cpp:143-168, :276-336): rows are dealt to numThreads sub-files
// (the reference's OpenMP dynamic schedule makes the assignment arbitrary; here row r goes to file r % numThreads).
struct SubFiles {
    std::string outputFile, folderPath;
    std::vector<std::string> dist_file_list, dist_index_list;
    std::vector<FILE *> fpArr, fpIndexArr;

    SubFiles(const std::string &out, int numThreads) : outputFile(out), folderPath(out + ".dir") {
        const std::string cmd = "mkdir -p " + folderPath;
        if (!system(cmd.c_str())) std::cerr << "success create: " << folderPath << std::endl;
        for (int i = 0; i < numThreads; i++) {
            dist_file_list.push_back(folderPath + '/' + outputFile + '.' + std::to_string(i));
            fpArr.push_back(fopen(dist_file_list.back().c_str(), "w+"));
            dist_index_list.push_back(outputFile + ".index." + std::to_string(i));
            fpIndexArr.push_back(fopen(dist_index_list.back().c_str(), "w+"));
            if (!fpArr.back() || !fpIndexArr.back()) {
                std::cerr << "ERROR: cannot open the distance sub-file: " << dist_file_list.back() << std::endl;
                exit(1);
            }
        }
    }

    void finish() {
        const int numThreads = (int)fpArr.size();
        for (int i = 0; i < numThreads; i++) {
            fclose(fpArr[i]);
            fclose(fpIndexArr[i]);
        }
        uint64_t totalSize = 0;
        uint64_t maxSize = 1LLU << 32;  // max total distance file size 4GB (dist.cpp:277)
        if (const char *lim = getenv("KSSD_B200_MERGE_LIMIT")) maxSize = strtoull(lim, nullptr, 10);  // tests: take the split branch without 4 GiB of text
        for (int i = 0; i < numThreads; i++) {
            struct stat st;
            st.st_size = 0;
            stat(dist_file_list[i].c_str(), &st);
            totalSize += st.st_size;
        }
        const int bufSize = 1 << 24;
        std::vector<char> buf(bufSize + 1);
        auto concat = [&](const std::string &from, FILE *to) {
            FILE *f = fopen(from.c_str(), "rb+");
            if (!f) return;
            while (true) {
                const size_t got = fread(buf.data(), 1, bufSize, f);
                fwrite(buf.data(), 1, got, to);
                if (got < (size_t)bufSize) break;
            }
            fclose(f);
        };
        if (totalSize <= maxSize) {
            FILE *com = fopen(outputFile.c_str(), "w");
            std::cerr << "-----save the output distance file: " << outputFile << std::endl;
            fprintf(com, " genome0\tgenome1\tcommon|size0|size1\tjaccard\tmashD\n");
            for (int i = 0; i < numThreads; i++) {
                concat(dist_file_list[i], com);
                remove(dist_file_list[i].c_str());
                remove(dist_index_list[i].c_str());
            }
            remove(folderPath.c_str());
            fclose(com);
        } else {
            std::cerr << "-----the output distance file is too big to merge into one single file, saving the result into directory: "
                      << folderPath << std::endl;
            const std::string outputIndexFile = outputFile + ".index";
            std::cerr << "-----save the index between genomes and distance sub-files into: " << outputIndexFile << std::endl;
            FILE *com = fopen(outputIndexFile.c_str(), "w+");
            fprintf(com, "genomeName\tdistFileName\n");
            for (int i = 0; i < numThreads; i++) {
                concat(dist_index_list[i], com);
                remove(dist_index_list[i].c_str());
            }
            fclose(com);
        }
    }
};

// Cells of one row in ascending column order, as the reference's j loop sees them.  The device reports only cells
// with common > 0 (kssd_b200_set_emit_zero_cells(ctx, 0)); when `dense` is set (a zero cell, distance exactly 1.0,
// passes the threshold) the unreported columns of [jbegin, jend) are visited too, with common = 0.
template <typename F>
inline void for_each_cell(const std::vector<kssd_hit> &hits, size_t kb, size_t ke, size_t jbegin, size_t jend, bool dense, F &&fn) {
    if (!dense) {
        for (size_t k = kb; k < ke; k++) fn((size_t)hits[k].col, (int)hits[k].common);
        return;
    }
    size_t k = kb;
    for (size_t j = jbegin; j < jend; j++) {
        int common = 0;
        if (k < ke && hits[k].col == j) common = (int)hits[k++].common;
        fn(j, common);
    }
}

// One context per GPU, every one holding the reference index (query rows are sharded, the index is
// replicated), and the query sketches uploaded once per GPU.
struct GpuPool {
    std::vector<kssd_ctx *> ctx;
    std::vector<const void *> d_qh;
    std::vector<const uint64_t *> d_qo;
    bool use64 = false;
    uint32_t n_query = 0;

    // index: what transSketches left on device 0 when it wrote exactly these files in this process (kssd_host::index_cache),
    // else the files (src/dist.cpp:83-130, :488-523).  Other GPUs: device 0's CSR copied over (32-bit hashes) or loaded
    // from the host arrays too.
    void setup(size_t want, bool use64_, const std::string &refSketchOut, const char *who, const std::vector<uint32_t> &ref_sizes, const Flat &queries) {
        use64 = use64_;
        n_query = (uint32_t)queries.sizes.size();
        const std::string indexFile = refSketchOut + ".index", dictFile = refSketchOut + ".dict";
        kssd_host::IndexCache &cache = kssd_host::index_cache();
        const bool cached = cache.valid && cache.use64 == use64 && cache.index_file == indexFile && cache.dict_file == dictFile &&
                            cache.n_genomes == ref_sizes.size();
        const size_t G = std::max<size_t>(1, std::min(want, kssd_host::device_count()));
        ctx = kssd_host::contexts_up_to(G);
        IndexFiles ix;
        if (cached) {
            std::cerr << "-----" << who << ": the index built by transSketches is still on the device, " << indexFile << " is not read back" << std::endl;
        } else {
            kssd_host::pending_writes().wait(&indexFile);
            read_index(indexFile, dictFile, who, use64, ix);
            load(ctx[0], ix, ref_sizes);
            cache = kssd_host::IndexCache{true, use64, indexFile, dictFile, ref_sizes.size()};
        }
        if (G > 1) {
            if (use64 && cached) {  // no device-to-device path for the 64-bit dictionary: the other GPUs read the files
                kssd_host::pending_writes().wait(&indexFile);
                read_index(indexFile, dictFile, who, use64, ix);
            }
            uint64_t n_uniq = 0, total = 0;
            uint32_t n_ref = 0;
            const uint32_t *d_uh = nullptr, *d_dict = nullptr, *d_rs = nullptr;
            const uint64_t *d_uo = nullptr;
            if (!use64) check(kssd_b200_index_device_views(ctx[0], &n_uniq, &total, &n_ref, &d_uh, &d_uo, &d_dict, &d_rs), "index_device_views");
#pragma omp parallel for num_threads((int)(G - 1)) schedule(static, 1)
            for (size_t g = 1; g < G; g++) {
                if (use64) load(ctx[g], ix, ref_sizes);
                else check(kssd_b200_index_adopt_device(ctx[g], n_uniq, total, n_ref, d_uh, d_uo, d_dict, d_rs), "index_adopt_device");
            }
        }
        d_qh.assign(G, nullptr);
        d_qo.assign(G, nullptr);
#pragma omp parallel for num_threads((int)G) schedule(static, 1)
        for (size_t g = 0; g < G; g++) {
            check(kssd_b200_set_emit_zero_cells(ctx[g], 0), "set_emit_zero_cells");
            check(kssd_b200_queries_upload(ctx[g], use64 ? (const void *)queries.h64.data() : (const void *)queries.h32.data(), queries.off.data(), n_query,
                                           use64 ? 8 : 4, &d_qh[g], &d_qo[g]), "queries_upload");
        }
    }

    void load(kssd_ctx *c, const IndexFiles &ix, const std::vector<uint32_t> &sizes) {
        if (use64)
            check(kssd_b200_index_load64(c, ix.hashSize, ix.hash64.data(), ix.counts.data(), ix.dict.data(), ix.totalIndex, sizes.data(),
                                         (uint32_t)sizes.size()), "index_load64");
        else
            check(kssd_b200_index_load(c, ix.hashSize, ix.counts.data(), ix.dict.data(), ix.totalIndex, sizes.data(), (uint32_t)sizes.size()),
                  "index_load");
    }

    void rows(size_t g, uint32_t row_begin, uint32_t row_end, int triangle, int kmer_size, double max_dist, int containment,
              std::vector<kssd_hit> *hits) const {
        if (use64)
            check(kssd_b200_dist_rows64(ctx[g], (const uint64_t *)d_qh[g], d_qo[g], n_query, KSSD_PTR_DEVICE, row_begin, row_end, triangle, kmer_size,
                                        max_dist, containment, collect_hits, hits), "dist_rows64");
        else
            check(kssd_b200_dist_rows(ctx[g], (const uint32_t *)d_qh[g], d_qo[g], n_query, KSSD_PTR_DEVICE, row_begin, row_end, triangle, kmer_size,
                                      max_dist, containment, collect_hits, hits), "dist_rows");
    }
};

// hits of the rows [r0, r1) sorted by (row, col); row_begin[r - r0] .. row_begin[r - r0 + 1] index the hits of row r
void group_by_row(std::vector<kssd_hit> &hits, size_t r0, size_t r1, std::vector<size_t> &row_begin) {
    std::sort(hits.begin(), hits.end(), [](const kssd_hit &a, const kssd_hit &b) { return a.row != b.row ? a.row < b.row : a.col < b.col; });
    row_begin.assign(r1 - r0 + 1, 0);
    for (const kssd_hit &h : hits) row_begin[h.row - r0 + 1]++;
    for (size_t r = 0; r < r1 - r0; r++) row_begin[r + 1] += row_begin[r];
}

// GPU g of G owns the sub-files [first, first + count) of the numThreads the reference would write
inline void files_of_gpu(size_t g, size_t G, int numThreads, int &first, int &count) {
    first = (int)(g * (size_t)numThreads / G);
    count = (int)((g + 1) * (size_t)numThreads / G) - first;
}

}  // namespace

void index_tridist(vector<sketch_t> &sketches, sketchInfo_t &info, string refSketchOut, string outputFile, int kmer_size,
                   double maxDist, int isContainment, int numThreads) {
    const double t0 = get_sec();
    const bool use64 = info.half_k - info.drlevel > 8;  // dist.cpp:29
    std::cerr << (use64 ? "-----use hash64 in index_tridist() " : "-----not use hash64 in index_tridist() ") << std::endl;
    if (numThreads < 1) numThreads = 1;
    const size_t numRef = sketches.size();
    Flat fl;
    flatten(sketches, use64, fl);
    const std::vector<uint32_t> &sizes = fl.sizes;
    GpuPool pool;
    // 100,000 sketches (5e9 cells) take 0.12 s on one B200: more GPUs only for jobs that are worth their start-up cost
    pool.setup(std::min<size_t>((size_t)numThreads, kssd_host::gpus_for((double)numRef * (double)numRef / 2.0, 2e10, "KSSD_B200_CELLS_PER_GPU")), use64,
               refSketchOut, "index_tridist()", sizes, fl);
    const size_t G = pool.ctx.size();
    const double t1 = get_sec();
    std::cerr << "===================time of read index and offset sketch file is: " << t1 - t0 << std::endl;

    SubFiles sub(outputFile, numThreads);
    std::cerr << "=====total: " << numRef << std::endl;
    const bool dense = 1.0 < maxDist;  // zero cells pass `mashD < maxDist` (dist.cpp:227-232)
    // dense: every unreported cell is printed as a zero cell, so the device must report EVERY cell with common > 0
    // (threshold +inf); the exact test below drops the ones whose distance does not pass
    const double devDist = dense ? INFINITY : maxDist;
    // query-row blocks per GPU, balanced on cells + per-row work; every GPU thread walks its block in pieces so that the hit
    // list stays bounded, formats them with its share of the threads and appends to its own sub-files as it goes
    const double avg_size = numRef ? (double)fl.off.back() / (double)numRef : 0.0;
    const std::vector<size_t> bounds = kssd_host::triangle_bounds(numRef, G, 40.0 * avg_size);
    const size_t piece = std::max<size_t>(64, std::min<size_t>(8192, (size_t)(256u << 20) / std::max<size_t>(numRef, 1)));
    omp_set_max_active_levels(2);
#pragma omp parallel for num_threads((int)G) schedule(static, 1)
    for (size_t g = 0; g < G; g++) {
        check(kssd_b200_set_max_neighbor(pool.ctx[g], 0), "set_max_neighbor");
        int f0, nf;
        files_of_gpu(g, G, numThreads, f0, nf);
        for (size_t p0 = bounds[g]; p0 < bounds[g + 1]; p0 += piece) {
            const size_t p1 = std::min(bounds[g + 1], p0 + piece);
            std::vector<kssd_hit> hits;
            pool.rows(g, (uint32_t)p0, (uint32_t)p1, 1, kmer_size, devDist, isContainment, &hits);
            std::vector<size_t> row_begin;
            group_by_row(hits, p0, p1, row_begin);
#pragma omp parallel for num_threads(nf) schedule(static, 1)
            for (int t = 0; t < nf; t++) {
                std::string strBuf;
                FILE *out = sub.fpArr[f0 + t];
                for (size_t i = p0 + ((size_t)t + (size_t)nf - p0 % (size_t)nf) % (size_t)nf; i < p1; i += (size_t)nf) {  // rows with i % nf == t
                    fprintf(sub.fpIndexArr[f0 + t], "%s\t%s\n", sketches[i].fileName.c_str(), sub.dist_file_list[f0 + t].c_str());
                    for_each_cell(hits, row_begin[i - p0], row_begin[i - p0 + 1], i + 1, numRef, dense, [&](size_t j, int common) {
                        const int size0 = (int)sizes[i], size1 = (int)sizes[j];
                        double jorc, d;
                        exact_distance(common, size0, size1, kmer_size, isContainment, jorc, d);
                        if (d < maxDist)  // strict, dist.cpp:232 / :251
                            strBuf += sketches[j].fileName + '\t' + sketches[i].fileName + '\t' + to_string(common) + '|' + to_string(size0) + '|' +
                                      to_string(size1) + '\t' + to_string(jorc) + '\t' + to_string(d) + '\n';
                        if (strBuf.size() > (1u << 22)) {
                            fwrite(strBuf.data(), 1, strBuf.size(), out);
                            strBuf.clear();
                        }
                    });
                }
                fwrite(strBuf.data(), 1, strBuf.size(), out);
            }
        }
    }
    const double t2 = get_sec();
    std::cerr << "===================time of multiple threads distance computing and save the subFile is: " << t2 - t1 << std::endl;
    sub.finish();
    std::cerr << "===================time of merge the subFiles into final files is: " << get_sec() - t2 << std::endl;
}

void index_dist(vector<sketch_t> &ref_sketches, sketchInfo_t &ref_info, string refSketchOut, vector<sketch_t> &query_sketches,
                string outputFile, int kmer_size, double maxDist, uint64_t maxNeighbor, bool isNeighbor, int isContainment, int numThreads) {
    const double t0 = get_sec();
    const bool use64 = ref_info.half_k - ref_info.drlevel > 8;  // dist.cpp:435
    std::cerr << (use64 ? "-----use hash64 in index_dist() " : "-----not use hash64 in index_dist() ") << std::endl;
    if (numThreads < 1) numThreads = 1;
    const size_t numQuery = query_sketches.size(), numRef = ref_sketches.size();
    Flat rfl, qfl;
    flatten(ref_sketches, use64, rfl);
    flatten(query_sketches, use64, qfl);
    const std::vector<uint32_t> &rsizes = rfl.sizes, &qsizes = qfl.sizes;
    GpuPool pool;
    pool.setup(std::min<size_t>((size_t)numThreads, kssd_host::gpus_for((double)numQuery * (double)numRef, 2e10, "KSSD_B200_CELLS_PER_GPU")), use64,
               refSketchOut, "index_dist()", rsizes, qfl);
    const size_t G = pool.ctx.size();
    const double t1 = get_sec();
    std::cerr << "===================time of read the index and dict sketch file is: " << t1 - t0 << std::endl;

    SubFiles sub(outputFile, numThreads);
    std::cerr << "=====total: " << numQuery << std::endl;
    const bool dense = 1.0 <= maxDist;  // zero cells pass `mashD <= maxDist` (dist.cpp:621-624)
    const double devDist = dense ? INFINITY : maxDist;  // dense: the device reports every cell with common > 0 (see index_tridist)
    // contiguous blocks of query rows per GPU, walked in pieces so that the hit list stays bounded when every cell passes
    const std::vector<size_t> bounds = kssd_host::rect_bounds(numQuery, G);
    const size_t piece = std::max<size_t>(1, std::min<size_t>(numQuery ? numQuery : 1, (size_t)(64u << 20) / std::max<size_t>(numRef, 1)));
    omp_set_max_active_levels(2);
#pragma omp parallel for num_threads((int)G) schedule(static, 1)
    for (size_t g = 0; g < G; g++) {
        // -N: the device selects the maxNeighbor smallest distances (+ ties) of every row among the cells with common > 0;
        // the bounded max-heap below (dist.cpp:625-640) runs over that superset
        check(kssd_b200_set_max_neighbor(pool.ctx[g], isNeighbor ? maxNeighbor : 0), "set_max_neighbor");
        int f0, nf;
        files_of_gpu(g, G, numThreads, f0, nf);
        for (size_t q0 = bounds[g]; q0 < bounds[g + 1]; q0 += piece) {
            const size_t q1 = std::min(bounds[g + 1], q0 + piece);
            std::vector<kssd_hit> hits;
            pool.rows(g, (uint32_t)q0, (uint32_t)q1, 0, kmer_size, devDist, isContainment, &hits);
            std::vector<size_t> row_begin;
            group_by_row(hits, q0, q1, row_begin);
#pragma omp parallel for num_threads(nf) schedule(static, 1)
            for (int t = 0; t < nf; t++) {
                std::string strBuf;
                FILE *out = sub.fpArr[f0 + t];
                for (size_t i = q0 + ((size_t)t + (size_t)nf - q0 % (size_t)nf) % (size_t)nf; i < q1; i += (size_t)nf) {  // rows with i % nf == t
                    fprintf(sub.fpIndexArr[f0 + t], "%s\t%s\n", query_sketches[i].fileName.c_str(), sub.dist_file_list[f0 + t].c_str());
                    const int size1 = (int)qsizes[i];
                    const size_t kb = row_begin[i - q0], ke = row_begin[i - q0 + 1];
                    priority_queue<DistInfo, vector<DistInfo>, cmpDistInfo> curQueue;  // dist.cpp:599
                    auto offer = [&](size_t j, int common) {  // one cell of the reference's j loop (dist.cpp:600-681)
                        const int size0 = (int)rsizes[j];
                        double jorc, d;
                        exact_distance(common, size0, size1, kmer_size, isContainment, jorc, d);
                        if (!(d <= maxDist)) return;  // inclusive, dist.cpp:624 / :661
                        if (isNeighbor) {               // bounded max-heap, dist.cpp:625-640
                            DistInfo di;
                            di.refName = ref_sketches[j].fileName;
                            di.common = common;
                            di.refSize = size0;
                            di.jorc = jorc;
                            di.dist = d;
                            if (curQueue.size() < maxNeighbor) curQueue.push(di);
                            else if (d < curQueue.top().dist) {
                                curQueue.push(di);
                                curQueue.pop();
                            }
                        } else {
                            strBuf += query_sketches[i].fileName + '\t' + ref_sketches[j].fileName + '\t' + to_string(common) + '|' + to_string(size0) +
                                      '|' + to_string(size1) + '\t' + to_string(jorc) + '\t' + to_string(d) + '\n';
                            if (strBuf.size() > (1u << 22)) {
                                fwrite(strBuf.data(), 1, strBuf.size(), out);
                                strBuf.clear();
                            }
                        }
                    };
                    if (!isNeighbor) {
                        for_each_cell(hits, kb, ke, 0, numRef, dense, offer);  // ascending ref index
                    } else if (maxNeighbor > 0) {
                        // The same heap operations as the reference's j loop, without walking all numRef columns: a zero cell
                        // (distance exactly 1.0) changes the heap only while it is not full or while its top is > 1.0, so the
                        // columns between two reported cells are visited only until then.
                        size_t j = 0;
                        for (size_t k = kb; k <= ke; k++) {
                            const size_t stop = k < ke ? (size_t)hits[k].col : numRef;
                            if (dense)
                                for (; j < stop; j++) {
                                    if (curQueue.size() >= maxNeighbor && !(1.0 < curQueue.top().dist)) break;
                                    offer(j, 0);
                                }
                            if (k < ke) offer(stop, (int)hits[k].common);
                            j = stop + 1;
                        }
                    }
                    while (!curQueue.empty()) {  // drained largest first, dist.cpp:683-689
                        const DistInfo di = curQueue.top();
                        strBuf += query_sketches[i].fileName + '\t' + di.refName + '\t' + to_string(di.common) + '|' + to_string(di.refSize) + '|' +
                                  to_string(size1) + '\t' + to_string(di.jorc) + '\t' + to_string(di.dist) + '\n';
                        curQueue.pop();
                    }
                    if (strBuf.size() > (1u << 22)) {
                        fwrite(strBuf.data(), 1, strBuf.size(), out);
                        strBuf.clear();
                    }
                }
                fwrite(strBuf.data(), 1, strBuf.size(), out);
            }
        }
        check(kssd_b200_set_max_neighbor(pool.ctx[g], 0), "set_max_neighbor");
    }
    const double t2 = get_sec();
    std::cerr << "===================time of multiple threads distance computing and save the subFile is: " << t2 - t1 << std::endl;
    sub.finish();
    std::cerr << "===================time of merge the subFiles into final files is: " << get_sec() - t2 << std::endl;
}
