/*
 * kssd_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, single-threaded CPU restatement of the two RabbitKSSD hot paths (sketch generation and
 * alldist/dist intersection counting + distances).  It exists only to check the CUDA path: tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it; nothing under
 * rabbitkssd_b200/ may import, link or call it (the product path has no CPU fallback).
 *
 * Parity pin: the reference ships no tests or golden vectors, so this restatement is
 * pinned against OUTPUTS OF THE REFERENCE BINARY ITSELF (oracle/build_ref.sh -> oracle/_ref/rabbit_kssd)
 * by tests/golden/make_golden.py; the committed fixtures under tests/golden/ are those outputs.
 *
 * Every function cites the reference file:line it follows (paths relative to the RabbitKSSD tree).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdio.h>

typedef struct {
    int half_k, half_subk, drlevel;
    int rev_add_move, half_outctx_len;
    int dim_start, dim_end;
    unsigned kmer_size;
    uint64_t domask, tupmask, undomask0, undomask1;
} oracle_params_t;

/* src/common.cpp:35-78 initParameter.  Returns 0 ok, -1 if the reference would abort (half_subk - drlevel < 3,
 * common.cpp:37-39) unless relax_guard != 0 (used only for the L5 configuration the reference cannot run). */
int kssd_oracle_init_params(int half_k, int half_subk, int drlevel, int relax_guard, oracle_params_t *p)
{
    if (half_subk - drlevel < 3 && !relax_guard) return -1;
    p->half_k = half_k;
    p->half_subk = half_subk;
    p->drlevel = drlevel;
    p->half_outctx_len = half_k - half_subk;
    p->rev_add_move = 4 * half_k - 2;
    p->kmer_size = 2 * half_k;
    p->dim_start = 0;
    p->dim_end = 1 << 4 * (half_subk - drlevel);
    int comp_bittl = 64 - 4 * half_k;
    uint64_t tupmask = 0xffffffffffffffffULL >> comp_bittl;
    uint64_t domask = (tupmask >> (4 * p->half_outctx_len)) << (2 * p->half_outctx_len);
    uint64_t undomask = (tupmask ^ domask) & tupmask;
    uint64_t undomask1 = undomask & (tupmask >> ((half_k + half_subk) * 2));
    uint64_t undomask0 = undomask ^ undomask1;
    p->domask = domask;
    p->tupmask = tupmask;
    p->undomask1 = undomask1;
    p->undomask0 = undomask0;
    return 0;
}

/* src/common.h:27-37 BaseMap: A/a 0, C/c 1, G/g 2, T/t 3, everything else -1.  The reference indexes the
 * 128-entry table with a signed char; bytes >= 0x80 are out of bounds there and are treated as invalid here. */
static inline int base_code(unsigned char ch)
{
    switch (ch) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return -1;
    }
}

/* src/shuffle.cpp:76-104 shuffleN + shuffle: identity, Fisher-Yates seeded 23, Fisher-Yates seeded `seed`,
 * both with glibc srand/rand and j = rand() % (i+1).  write_shuffle_dim_file (shuffle.cpp:48-54) uses
 * seed = id = (k<<8)+(subk<<4)+drlevel.  out must hold 16^half_subk ints. */
void kssd_oracle_shuffle_dim(int half_k, int half_subk, int drlevel, int32_t *out)
{
    int n = 1 << 4 * half_subk;
    for (int i = 0; i < n; i++) out[i] = i;
    unsigned seeds[2] = {23u, (unsigned)((half_k << 8) + (half_subk << 4) + drlevel)};
    for (int s = 0; s < 2; s++) {
        srand(seeds[s]);
        for (int i = n - 1; i > 0; i--) {
            int j = rand() % (i + 1);
            int32_t tmp = out[i];
            out[i] = out[j];
            out[j] = tmp;
        }
    }
}

/* Scan ONE parsed sequence record and append every emitted hash (duplicates included) to out.
 * Follows src/sketch.cpp:487-532 (FASTA small-file loop; identical arithmetic in consumer_fasta_task :196-231)
 * and, when qual != NULL, the FASTQ validity rule quality[i] >= leastQual of :280 / :786 (raw ASCII compare
 * on a signed char).  The reference probes a hash map of the retained subset (sketch.cpp:338-345); probing
 * the dense permutation with the same predicate (value in [dim_start, dim_end)) is equivalent.
 * Returns number of hashes appended; stops appending (but keeps counting) past cap. */
uint64_t kssd_oracle_scan_record(const oracle_params_t *p, const int32_t *shuffled_dim,
                                 const unsigned char *seq, uint64_t length,
                                 const signed char *qual, int least_qual,
                                 uint64_t *out, uint64_t cap)
{
    uint64_t n = 0;
    uint64_t tuple = 0, rvs_tuple = 0, uni_tuple, dr_tuple, pfilter;
    unsigned base = 1;
    for (uint64_t i = 0; i < length; i++) {
        int basenum = base_code(seq[i]);
        if (basenum != -1 && (qual == NULL || (int)qual[i] >= least_qual)) {
            tuple = ((tuple << 2) | (uint64_t)basenum) & p->tupmask;
            rvs_tuple = (rvs_tuple >> 2) + (((uint64_t)basenum ^ 3ULL) << p->rev_add_move);
            base++;
        } else {
            base = 1;
        }
        if (base > p->kmer_size) {
            uni_tuple = tuple < rvs_tuple ? tuple : rvs_tuple;
            int dim_id = (int)((uni_tuple & p->domask) >> (p->half_outctx_len * 2));
            int32_t v = shuffled_dim[dim_id];
            if (v >= p->dim_end || v < p->dim_start) continue;
            pfilter = (uint64_t)(v - p->dim_start);
            dr_tuple = (((uni_tuple & p->undomask0) |
                         ((uni_tuple & p->undomask1) << (p->kmer_size * 2 - p->half_outctx_len * 4)))
                        >> (p->drlevel * 4)) | pfilter;
            if (n < cap) out[n] = dr_tuple;
            n++;
        }
    }
    return n;
}

static int cmp_u64(const void *a, const void *b)
{
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}

/* Per-genome set semantics: FASTA = unordered_set insert (sketch.cpp:526-529, :537-550); FASTQ = count per
 * hash over the whole file, keep count >= leastNumKmer (sketch.cpp:300-307, :832-849).  The reference's
 * iteration order is unspecified; this returns the sorted representative.  In place; returns kept count.
 * With use64 == 0 hashes are first truncated to 32 bits as the uint32_t containers do. */
uint64_t kssd_oracle_finish_genome(uint64_t *hashes, uint64_t n, int least_num_kmer, int use64)
{
    if (!use64) for (uint64_t i = 0; i < n; i++) hashes[i] = (uint32_t)hashes[i];
    qsort(hashes, n, sizeof(uint64_t), cmp_u64);
    uint64_t w = 0, i = 0;
    while (i < n) {
        uint64_t j = i;
        while (j < n && hashes[j] == hashes[i]) j++;
        if ((int64_t)(j - i) >= (int64_t)least_num_kmer) hashes[w++] = hashes[i];
        i = j;
    }
    return w;
}

/* src/sketch.cpp:970-1017 transSketches, 32-bit path: invert genome -> hashes into hash -> ascending genome
 * ids.  Output is the CONTENT of the two files: dict = posting lists concatenated in ascending hash order
 * (.dict), and for the dense form counts[hash] (.index payload).  Here the sparse form is returned:
 * uniq_hash[u], uniq_count[u] in ascending hash order; dict[total].  Inputs: hashes of all genomes
 * concatenated, genome g owning [genome_off[g], genome_off[g+1]); hashes unique within a genome.
 * Returns number of distinct hashes. */
typedef struct { uint32_t hash, genome; } pair_t;
static int cmp_pair(const void *a, const void *b)
{
    const pair_t *x = a, *y = b;
    if (x->hash != y->hash) return x->hash < y->hash ? -1 : 1;
    return x->genome < y->genome ? -1 : x->genome > y->genome;
}
uint64_t kssd_oracle_index_build(const uint32_t *hashes, const uint64_t *genome_off, uint32_t n_genomes,
                                 uint32_t *dict, uint32_t *uniq_hash, uint32_t *uniq_count)
{
    uint64_t total = genome_off[n_genomes];
    pair_t *pairs = malloc((total ? total : 1) * sizeof(pair_t));
    for (uint32_t g = 0; g < n_genomes; g++)
        for (uint64_t i = genome_off[g]; i < genome_off[g + 1]; i++) {
            pairs[i].hash = hashes[i];
            pairs[i].genome = g;
        }
    qsort(pairs, total, sizeof(pair_t), cmp_pair);
    uint64_t u = 0;
    for (uint64_t i = 0; i < total; i++) {
        dict[i] = pairs[i].genome;
        if (i == 0 || pairs[i].hash != pairs[i - 1].hash) {
            uniq_hash[u] = pairs[i].hash;
            uniq_count[u] = 0;
            u++;
        }
        uniq_count[u - 1]++;
    }
    free(pairs);
    return u;
}

/* src/dist.cpp:193-204 / :577-588: dense per-query counter row.  counts[n_ref] is zeroed then, for every
 * hash of the query, every posting id is incremented.  Sparse index: binary search replaces the dense
 * offset[] lookup of dist.cpp:197-198 (same posting range). */
void kssd_oracle_count_row(const uint32_t *q_hashes, uint64_t q_n,
                           const uint32_t *uniq_hash, const uint64_t *uniq_off, uint64_t n_uniq,
                           const uint32_t *dict, uint32_t n_ref, int32_t *counts)
{
    memset(counts, 0, (size_t)n_ref * sizeof(int32_t));
    for (uint64_t j = 0; j < q_n; j++) {
        uint32_t h = q_hashes[j];
        uint64_t lo = 0, hi = n_uniq;
        while (lo < hi) {
            uint64_t mid = (lo + hi) >> 1;
            if (uniq_hash[mid] < h) lo = mid + 1; else hi = mid;
        }
        if (lo == n_uniq || uniq_hash[lo] != h) continue;
        for (uint64_t k = uniq_off[lo]; k < uniq_off[lo + 1]; k++) counts[dict[k]]++;
    }
}

/* src/dist.cpp:216-232 (Jaccard -> Mash distance) and :237-251 (containment -> AAF distance); identical
 * expressions at :607-624 and :648-661.  jorc receives the Jaccard / containment value. */
double kssd_oracle_distance(int common, int size0, int size1, int kmer_size, int is_containment, double *jorc)
{
    if (!is_containment) {
        int denom = size0 + size1 - common;
        double jaccard;
        if (size0 == 0 || size1 == 0) jaccard = 0.0;
        else jaccard = (double)common / denom;
        double mashD;
        if (jaccard == 1.0) mashD = 0.0;
        else if (jaccard == 0.0) mashD = 1.0;
        else mashD = (double)-1.0 / kmer_size * log((2 * jaccard) / (1.0 + jaccard));
        *jorc = jaccard;
        return mashD;
    } else {
        int denom = size0 < size1 ? size0 : size1;
        double containment;
        if (size0 == 0 || size1 == 0) containment = 0.0;
        else containment = (double)common / denom;
        double AafD;
        if (containment == 1.0) AafD = 0.0;
        else if (containment == 0.0) AafD = 1.0;
        else AafD = (double)-1.0 / kmer_size * log(containment);
        *jorc = containment;
        return AafD;
    }
}

/* One row of alldist (dist.cpp:207-255): for j in (i, n): keep iff dist < maxDist (strict, :232/:251).
 * One row of dist (dist.cpp:600-682, non-neighbor mode): for j in [0, n_ref): keep iff dist <= maxDist
 * (:624/:661), size0 = ref size, size1 = query size.  hits: (col, common) pairs; returns count. */
uint64_t kssd_oracle_row_hits(const int32_t *counts, uint32_t j_begin, uint32_t n_ref,
                              const uint32_t *ref_sizes, uint32_t row_size, int row_is_size0,
                              int kmer_size, double max_dist, int inclusive, int is_containment,
                              uint32_t *hit_col, uint32_t *hit_common, double *hit_jorc, double *hit_dist,
                              uint64_t cap)
{
    uint64_t n = 0;
    for (uint32_t j = j_begin; j < n_ref; j++) {
        int common = counts[j];
        int size0 = row_is_size0 ? (int)row_size : (int)ref_sizes[j];
        int size1 = row_is_size0 ? (int)ref_sizes[j] : (int)row_size;
        double jorc;
        double d = kssd_oracle_distance(common, size0, size1, kmer_size, is_containment, &jorc);
        int keep = inclusive ? (d <= max_dist) : (d < max_dist);
        if (keep) {
            if (n < cap) {
                hit_col[n] = j;
                hit_common[n] = (uint32_t)common;
                if (hit_jorc) hit_jorc[n] = jorc;
                if (hit_dist) hit_dist[n] = d;
            }
            n++;
        }
    }
    return n;
}

/* Multithreaded variants used only as the timed CPU baseline ("port") when oracle/_ref is unavailable:
 * the same per-genome / per-row functions above under an OpenMP parallel-for over genomes (the reference's
 * own strategy, sketch.cpp:455, dist.cpp:174).  Compiled in only with -fopenmp. */
#ifdef _OPENMP
#include <omp.h>
int kssd_oracle_max_threads(void) { return omp_get_max_threads(); }
#else
int kssd_oracle_max_threads(void) { return 1; }
#endif

/* Sketch a batch: genome g = records [rec_first[g], rec_first[g+1]) ; record r = seq[rec_off[r], rec_off[r+1]).
 * out_hashes must hold out_cap u64 per genome (genome g writes at g*out_cap); out_n[g] = kept count
 * (sorted, unique / count-filtered). */
void kssd_oracle_sketch_batch(const oracle_params_t *p, const int32_t *shuffled_dim,
                              const unsigned char *seq, const signed char *qual,
                              const uint64_t *rec_off, const uint64_t *rec_first, uint32_t n_genomes,
                              int least_qual, int least_num_kmer, int use64,
                              uint64_t *out_hashes, uint64_t out_cap, uint64_t *out_n)
{
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic)
#endif
    for (int64_t g = 0; g < (int64_t)n_genomes; g++) {
        uint64_t *dst = out_hashes + (uint64_t)g * out_cap;
        uint64_t n = 0;
        for (uint64_t r = rec_first[g]; r < rec_first[g + 1]; r++) {
            uint64_t got = kssd_oracle_scan_record(p, shuffled_dim, seq + rec_off[r], rec_off[r + 1] - rec_off[r],
                                                   qual ? qual + rec_off[r] : NULL, least_qual,
                                                   dst + (n < out_cap ? n : out_cap), n < out_cap ? out_cap - n : 0);
            n += got;
        }
        if (n > out_cap) { out_n[g] = UINT64_MAX; continue; }
        out_n[g] = kssd_oracle_finish_genome(dst, n, least_num_kmer, use64);
    }
}

/* alldist distance loop over rows [row_begin, row_end) (dist.cpp:174-258): returns number of kept pairs and
 * (optionally) writes (row, col, common) triples. */
uint64_t kssd_oracle_alldist(const uint32_t *hashes, const uint64_t *genome_off, uint32_t n,
                             const uint32_t *uniq_hash, const uint64_t *uniq_off, uint64_t n_uniq,
                             const uint32_t *dict, uint32_t row_begin, uint32_t row_end,
                             int kmer_size, double max_dist, int is_containment,
                             uint32_t *hit_row, uint32_t *hit_col, uint32_t *hit_common, uint64_t cap)
{
    uint32_t *sizes = malloc((size_t)(n ? n : 1) * sizeof(uint32_t));
    for (uint32_t g = 0; g < n; g++) sizes[g] = (uint32_t)(genome_off[g + 1] - genome_off[g]);
    uint64_t total = 0;
#ifdef _OPENMP
#pragma omp parallel
#endif
    {
        int32_t *counts = malloc((size_t)(n ? n : 1) * sizeof(int32_t));
        uint32_t *cols = malloc((size_t)(n ? n : 1) * sizeof(uint32_t));
        uint32_t *coms = malloc((size_t)(n ? n : 1) * sizeof(uint32_t));
#ifdef _OPENMP
#pragma omp for schedule(dynamic)
#endif
        for (int64_t i = row_begin; i < (int64_t)row_end; i++) {
            kssd_oracle_count_row(hashes + genome_off[i], genome_off[i + 1] - genome_off[i],
                                  uniq_hash, uniq_off, n_uniq, dict, n, counts);
            uint64_t k = kssd_oracle_row_hits(counts, (uint32_t)i + 1, n, sizes, sizes[i], 1, kmer_size,
                                              max_dist, 0, is_containment, cols, coms, NULL, NULL, n);
            uint64_t at;
#ifdef _OPENMP
#pragma omp critical
#endif
            { at = total; total += k; }
            if (hit_row)
                for (uint64_t t = 0; t < k && at + t < cap; t++) {
                    hit_row[at + t] = (uint32_t)i;
                    hit_col[at + t] = cols[t];
                    hit_common[at + t] = coms[t];
                }
        }
        free(counts); free(cols); free(coms);
    }
    free(sizes);
    return total;
}

/* dist (ref x query) distance loop over query rows [row_begin,row_end) (dist.cpp:560-692, non-neighbor mode). */
uint64_t kssd_oracle_dist(const uint32_t *q_hashes, const uint64_t *q_off, uint32_t row_begin, uint32_t row_end,
                          const uint32_t *ref_sizes, uint32_t n_ref,
                          const uint32_t *uniq_hash, const uint64_t *uniq_off, uint64_t n_uniq,
                          const uint32_t *dict, int kmer_size, double max_dist, int is_containment,
                          uint32_t *hit_row, uint32_t *hit_col, uint32_t *hit_common, uint64_t cap)
{
    uint64_t total = 0;
#ifdef _OPENMP
#pragma omp parallel
#endif
    {
        int32_t *counts = malloc((size_t)(n_ref ? n_ref : 1) * sizeof(int32_t));
        uint32_t *cols = malloc((size_t)(n_ref ? n_ref : 1) * sizeof(uint32_t));
        uint32_t *coms = malloc((size_t)(n_ref ? n_ref : 1) * sizeof(uint32_t));
#ifdef _OPENMP
#pragma omp for schedule(dynamic)
#endif
        for (int64_t i = row_begin; i < (int64_t)row_end; i++) {
            uint32_t qs = (uint32_t)(q_off[i + 1] - q_off[i]);
            kssd_oracle_count_row(q_hashes + q_off[i], qs, uniq_hash, uniq_off, n_uniq, dict, n_ref, counts);
            uint64_t k = kssd_oracle_row_hits(counts, 0, n_ref, ref_sizes, qs, 0, kmer_size,
                                              max_dist, 1, is_containment, cols, coms, NULL, NULL, n_ref);
            uint64_t at;
#ifdef _OPENMP
#pragma omp critical
#endif
            { at = total; total += k; }
            if (hit_row)
                for (uint64_t t = 0; t < k && at + t < cap; t++) {
                    hit_row[at + t] = (uint32_t)i;
                    hit_col[at + t] = cols[t];
                    hit_common[at + t] = coms[t];
                }
        }
        free(counts); free(cols); free(coms);
    }
    return total;
}

/* ---- synthetic inputs for the CPU legs of bench.py (not part of the restated algorithm) -------------------------
 * The same genomes rabbitkssd_b200/synth.py:family() and synth.family_device() generate (counter-based splitmix64:
 * base genome = 2 bits per base from the stream of `seed`; member m = base with i.i.d. substitutions at rate
 * max_div * m / (n_members - 1) drawn from the stream of seed * 1000003 + m), written as FASTA files so that the
 * reference binary can be timed on exactly the workload the GPU arm runs. */
static inline uint64_t splitmix_at(uint64_t seed, uint64_t counter)
{
    uint64_t z = counter * 0x9E3779B97F4A7C15ULL + seed;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

void kssd_oracle_synth_genome(uint64_t seed, int member, int n_members, uint64_t length, double max_div, uint8_t *out)
{
    static const char acgt[4] = {'A', 'C', 'G', 'T'};
    const uint64_t seed2 = seed * 1000003ULL + (uint64_t)member;
    const double rate = member ? max_div * (double)member / (double)(n_members > 1 ? n_members - 1 : 1) : 0.0;
    const uint64_t thresh = (uint64_t)(rate * 9007199254740992.0); /* int(rate * 2.0 ** 53) */
    for (uint64_t w = 0; w * 32 < length; w++) {
        const uint64_t r = splitmix_at(seed, w + 1);
        const uint64_t end = (w + 1) * 32 < length ? (w + 1) * 32 : length;
        for (uint64_t i = w * 32; i < end; i++) {
            unsigned code = (unsigned)(r >> (2 * (i - w * 32))) & 3u;
            if (member) {
                const uint64_t m = splitmix_at(seed2, i + 1);
                if ((m >> 11) < thresh) code = (code + (unsigned)((m & 0x7FF) % 3) + 1u) & 3u;
            }
            out[i] = (uint8_t)acgt[code];
        }
    }
}

/* Writes genomes [first, first + n) of the family series (family f = index / n_members has seed seed0 + 101 * f) as
 * <dir>/g<index>.fna: n_contigs records "<name>_<c> synthetic", 80-column lines (synth.write_fasta).  Returns the
 * number of files written, < n on an I/O error. */
int kssd_oracle_synth_fasta_files(const char *dir, uint64_t seed0, uint32_t first, uint32_t n, int n_members, uint64_t length,
                                  double max_div, int n_contigs)
{
    int written = 0;
#pragma omp parallel for schedule(dynamic) reduction(+ : written)
    for (uint32_t q = 0; q < n; q++) {
        const uint32_t g = first + q;
        uint8_t *buf = (uint8_t *)malloc(length + 1);
        char path[4096];
        snprintf(path, sizeof(path), "%s/g%u.fna", dir, g);
        FILE *f = buf ? fopen(path, "wb") : NULL;
        if (!f) { free(buf); continue; }
        kssd_oracle_synth_genome(seed0 + 101ULL * (g / (uint32_t)n_members), (int)(g % (uint32_t)n_members), n_members, length, max_div, buf);
        char *line = (char *)malloc((size_t)(length + length / 80 + 2));
        int ok = line != NULL;
        for (int c = 0; ok && c < n_contigs; c++) {
            const uint64_t b0 = length * (uint64_t)c / (uint64_t)n_contigs, b1 = length * (uint64_t)(c + 1) / (uint64_t)n_contigs;
            fprintf(f, ">g%u.fna_%d synthetic\n", g, c);
            size_t o = 0;
            for (uint64_t i = b0; i < b1; i += 80) {
                const uint64_t e = i + 80 < b1 ? i + 80 : b1;
                memcpy(line + o, buf + i, e - i);
                o += e - i;
                line[o++] = '\n';
            }
            ok = fwrite(line, 1, o, f) == o;
        }
        free(line);
        ok = (fclose(f) == 0) && ok;
        free(buf);
        written += ok;
    }
    return written;
}
