// Test harness (CPU): prints the pieces RawSlicer (rabbitkssd_b200/csrc/host/raw_slicer.h) cuts a file into.
#include <cstdio>
#include <cstdlib>
#include <string>

#include "raw_slicer.h"

int main(int argc, char **argv) {
    if (argc < 3) return 2;
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 2;
    std::string t;
    char buf[65536];
    size_t r;
    while ((r = fread(buf, 1, sizeof buf, f)) > 0) t.append(buf, r);
    fclose(f);
    const size_t piece = (size_t)atoll(argv[2]);
    if (argc > 3 && std::string(argv[3]) == "fastq") {  // the cuts of the raw FASTQ reader: "<begin> <end of the whole records>" per piece
        size_t pos = 0;
        while (pos < t.size()) {
            const size_t n = std::min(t.size() - pos, piece);
            const size_t whole = kssd_host::fastq_whole_records((const unsigned char *)t.data() + pos, n);
            printf("%zu %zu\n", pos, pos + whole);
            if (whole == 0 || pos + n == t.size()) break;
            pos += whole;
        }
        return 0;
    }
    kssd_host::RawSlicer sl;
    size_t pos = 0;
    const size_t n = t.size();
    while (pos < n) {
        size_t end = std::min(n, pos + piece);
        if (end < n && t[end - 1] == '\r' && end - 1 > pos) end--;
        unsigned char prefix[kssd_host::RawSlicer::kCtx + 1];
        size_t plen = 0;
        const int flags = pos == 0 ? 0 : sl.next_prefix(prefix, plen);
        printf("%zu %zu %d %zu ", pos, end, flags, plen);
        for (size_t i = 0; i < plen; i++) printf("%02x", prefix[i]);
        printf("\n");
        if (end < n) sl.advance((const unsigned char *)t.data() + pos, end - pos, (unsigned char)t[end]);
        pos = end;
    }
    return 0;
}
