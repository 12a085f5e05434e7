// Binary CSR file, basic format of the reference (apps/output_cpp/gm_graph/src/gm_graph_binary_loader.cc:19-26):
//   big-endian [magic 0x03939999][sizeof node_t][sizeof edge_t][N][M][begin[N+1]][node_idx[M]]
// The reference reads/writes one element per fread/fwrite (:140-178, :243-251); here the arrays move in
// bulk and are byte-swapped in place.  Only 4-byte node/edge files (the default build) are accepted.
#include <arpa/inet.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
#include "gmc_common.cuh"

using namespace gmc;

static const uint32_t MAGIC_WORD_BIN = 0x03939999u;

static int read_header(FILE* f, const char* path, int32_t* n, int32_t* m, bool* flipped) {
    uint32_t h[5];
    GMC_REQUIRE(fread(h, 4, 5, f) == 5, GMC_ERR_IO, "%s: wrong file format (short header)", path);
    if (ntohl(h[0]) == MAGIC_WORD_BIN) *flipped = false;
    else if (h[0] == MAGIC_WORD_BIN) *flipped = true;     // "old flipped format", loader :78-80
    else { set_error("%s: wrong file format, KEY mismatch: %08x, expected %08x", path, ntohl(h[0]), MAGIC_WORD_BIN); return GMC_ERR_IO; }
    uint32_t ns = *flipped ? h[1] : ntohl(h[1]), es = *flipped ? h[2] : ntohl(h[2]);
    GMC_REQUIRE(ns == 4 && es == 4, GMC_ERR_IO, "%s: node_t/edge_t size %u/%u not supported by this build (4/4)", path, ns, es);
    *n = (int32_t)(*flipped ? h[3] : ntohl(h[3]));
    *m = (int32_t)(*flipped ? h[4] : ntohl(h[4]));
    GMC_REQUIRE(*n >= 0 && *m >= 0, GMC_ERR_IO, "%s: negative sizes", path);
    return GMC_OK;
}

static void swap_words(int32_t* p, size_t count) {
    #pragma omp parallel for
    for (size_t i = 0; i < count; i++) p[i] = (int32_t)__builtin_bswap32((uint32_t)p[i]);
}

extern "C" {

int gmc_file_read_header(const char* path, int32_t* n, int32_t* m) {
    GMC_REQUIRE(path && n && m, GMC_ERR_ARG, "gmc_file_read_header: NULL");
    FILE* f = fopen(path, "rb");
    GMC_REQUIRE(f, GMC_ERR_IO, "cannot open %s for reading", path);
    bool flipped;
    int rc = read_header(f, path, n, m, &flipped);
    fclose(f);
    return rc;
}

int gmc_file_read(const char* path, int32_t* begin, int32_t* node_idx) {
    GMC_REQUIRE(path && begin, GMC_ERR_ARG, "gmc_file_read: NULL");
    FILE* f = fopen(path, "rb");
    GMC_REQUIRE(f, GMC_ERR_IO, "cannot open %s for reading", path);
    int32_t n, m; bool flipped;
    int rc = read_header(f, path, &n, &m, &flipped);
    if (rc == GMC_OK && fread(begin, 4, (size_t)n + 1, f) != (size_t)n + 1) { set_error("%s: error reading node begin array", path); rc = GMC_ERR_IO; }
    if (rc == GMC_OK && m > 0 && (!node_idx || fread(node_idx, 4, (size_t)m, f) != (size_t)m)) { set_error("%s: error reading edge-end array", path); rc = GMC_ERR_IO; }
    fclose(f);
    if (rc == GMC_OK && !flipped) { swap_words(begin, (size_t)n + 1); if (m) swap_words(node_idx, (size_t)m); }
    // a truncated or corrupt file must be an error here, not an out-of-bounds access in a kernel: offsets start at 0, never
    // decrease, end at M; every neighbour id is a vertex
    if (rc == GMC_OK) {
        long bad = 0;
        if (begin[0] != 0 || begin[n] != m) bad = 1;
#pragma omp parallel for reduction(+ : bad) schedule(static)
        for (int64_t v = 0; v < n; v++) bad += begin[v + 1] < begin[v];
#pragma omp parallel for reduction(+ : bad) schedule(static)
        for (int64_t e2 = 0; e2 < m; e2++) bad += (node_idx[e2] < 0) | (node_idx[e2] >= n);
        if (bad) { set_error("%s: not a valid CSR (offsets must start at 0, be non-decreasing and end at %d; neighbour ids must be below %d)", path, m, n); rc = GMC_ERR_IO; }
    }
    return rc;
}

int gmc_file_write(const char* path, int32_t n, int32_t m, const int32_t* begin, const int32_t* node_idx) {
    GMC_REQUIRE(path && begin && (node_idx || m == 0) && n >= 0 && m >= 0, GMC_ERR_ARG, "gmc_file_write: bad args");
    FILE* f = fopen(path, "wb");
    GMC_REQUIRE(f, GMC_ERR_IO, "cannot open %s for writing", path);
    uint32_t h[5] = {htonl(MAGIC_WORD_BIN), htonl(4), htonl(4), htonl((uint32_t)n), htonl((uint32_t)m)};
    bool ok = fwrite(h, 4, 5, f) == 5;
    const size_t CH = 1 << 20;
    std::vector<uint32_t> buf(CH);
    auto put = [&](const int32_t* p, size_t count) {
        for (size_t o = 0; ok && o < count; o += CH) {
            size_t c = count - o < CH ? count - o : CH;
            for (size_t i = 0; i < c; i++) buf[i] = __builtin_bswap32((uint32_t)p[o + i]);
            ok = fwrite(buf.data(), 4, c, f) == c;
        }
    };
    put(begin, (size_t)n + 1);
    put(node_idx, (size_t)m);
    ok = (fclose(f) == 0) && ok;
    GMC_REQUIRE(ok, GMC_ERR_IO, "%s: write failed", path);
    return GMC_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// EDGE_LIST text ingest: gm_edge_list_graph_reader::builtGraph (apps/output_cpp/gm_graph/src/
// gm_edge_list_graph_reader.cc:68-171).  Lines "src dst" or "src *" (vertex only), blank lines and lines starting
// with '#' skipped (:99,:140), tokens separated by blanks / tabs (:101,:105); every id that occurs becomes a vertex,
// numbered densely in ascending raw-id order (:112-117, ids below LOOKUP_SIZE = 80 000 000, :66).  The reference reads
// the file three times with getline/strtok and inserts into a std::map; here the file is mapped once, the lines
// are parsed by all host threads, and the (src, dst) pairs go to the device as 64-bit keys where one radix sort
// yields the semi-sorted CSR (gmc_gen.cu: gmc_graph_from_edge_keys).  Input need not be sorted by source
// (the reference asserts it, :146); on sorted input the result is identical.
// ---------------------------------------------------------------------------------------------------------
#include <fcntl.h>
#include <omp.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

static const int64_t EDGE_LIST_LOOKUP_SIZE = 80000000;

namespace {
struct ParsedLines {
    std::vector<int32_t> src, dst, lone;
    int64_t max_id = -1;
    long bad_line = -1;          // byte offset of the first malformed line
};
inline bool is_blank(char c) { return c == ' ' || c == '\t'; }
// one token of decimal digits -> id; false when it is not a non-negative integer below LOOKUP_SIZE
inline bool parse_id(const char*& p, const char* end, int64_t& v) {
    const char* q = p;
    if (q < end && *q == '+') q++;
    if (q >= end || *q < '0' || *q > '9') return false;
    int64_t x = 0;
    while (q < end && *q >= '0' && *q <= '9') { x = x * 10 + (*q - '0'); if (x >= EDGE_LIST_LOOKUP_SIZE) return false; q++; }
    if (q < end && !is_blank(*q) && *q != '\r' && *q != '\n') return false;
    p = q; v = x;
    return true;
}
void parse_range(const char* base, const char* p, const char* end, ParsedLines& out) {
    while (p < end) {
        const char* line = p;
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        p = eol + 1;
        if (eol == line || *line == '#') continue;
        const char* q = line;
        while (q < eol && is_blank(*q)) q++;
        if (q == eol || *q == '\r') continue;
        int64_t a, b;
        bool ok = parse_id(q, eol, a);
        while (ok && q < eol && is_blank(*q)) q++;
        if (ok && q < eol && *q == '*') {
            out.lone.push_back((int32_t)a);
            if (a > out.max_id) out.max_id = a;
            continue;
        }
        ok = ok && parse_id(q, eol, b);
        if (!ok) { if (out.bad_line < 0) out.bad_line = (long)(line - base); continue; }
        out.src.push_back((int32_t)a); out.dst.push_back((int32_t)b);
        if (a > out.max_id) out.max_id = a;
        if (b > out.max_id) out.max_id = b;
    }
}
}  // namespace

// internal (gmc_gen.cu): edge keys (src << 32 | dst) in host memory -> device graph, semi-sorted (+ reverse CSR)
extern "C" int gmc_graph_from_edge_keys(gmc_graph** out, int32_t n, int32_t m, const uint64_t* keys_host, int make_reverse);

extern "C" int gmc_edge_list_read(const char* path, gmc_graph** out, int make_reverse) {
    GMC_REQUIRE(path && out, GMC_ERR_ARG, "gmc_edge_list_read: NULL");
    *out = nullptr;
    int fd = open(path, O_RDONLY);
    GMC_REQUIRE(fd >= 0, GMC_ERR_IO, "gmc_edge_list_read: cannot open %s", path);
    struct stat sb;
    if (fstat(fd, &sb) != 0) { close(fd); set_error("gmc_edge_list_read: cannot stat %s", path); return GMC_ERR_IO; }
    const size_t sz = (size_t)sb.st_size;
    const char* base = sz ? (const char*)mmap(nullptr, sz, PROT_READ, MAP_PRIVATE, fd, 0) : "";
    close(fd);
    if (sz && base == MAP_FAILED) { set_error("gmc_edge_list_read: cannot map %s", path); return GMC_ERR_IO; }
    const int T = sz > (1u << 20) ? omp_get_max_threads() : 1;
    std::vector<ParsedLines> parts((size_t)T);
    std::vector<size_t> cut((size_t)T + 1);
    for (int t = 0; t <= T; t++) {          // chunk borders moved forward to the next line start
        size_t c = sz / (size_t)T * (size_t)t;
        if (t == T) c = sz;
        else if (t > 0) { const char* nl = (const char*)memchr(base + c, '\n', sz - c); c = nl ? (size_t)(nl - base) + 1 : sz; }
        cut[t] = c;
    }
#pragma omp parallel for num_threads(T) schedule(static, 1)
    for (int t = 0; t < T; t++)
        if (cut[t + 1] > cut[t]) parse_range(base, base + cut[t], base + cut[t + 1], parts[t]);
    if (sz) munmap((void*)base, sz);
    int64_t max_id = -1, m64 = 0;
    for (auto& p : parts) {
        if (p.bad_line >= 0) { set_error("gmc_edge_list_read: %s: malformed line at byte %ld (want \"src dst\" or \"src *\", ids in [0, %lld))", path, p.bad_line, (long long)EDGE_LIST_LOOKUP_SIZE); return GMC_ERR_IO; }
        if (p.max_id > max_id) max_id = p.max_id;
        m64 += (int64_t)p.src.size();
    }
    GMC_REQUIRE(m64 <= 0x7fffff00LL, GMC_ERR_ARG, "gmc_edge_list_read: %lld edges exceed the 32-bit edge_t build", (long long)m64);
    // ids that occur, numbered densely in ascending raw order
    std::vector<int32_t> lookup((size_t)(max_id + 1), 0);
#pragma omp parallel for num_threads(T) schedule(static, 1)
    for (int t = 0; t < T; t++) {
        for (int32_t a : parts[t].src) lookup[a] = 1;       // racing writers all store 1
        for (int32_t b : parts[t].dst) lookup[b] = 1;
        for (int32_t a : parts[t].lone) lookup[a] = 1;
    }
    int32_t n = 0;
    for (int64_t i = 0; i <= max_id; i++) { const int32_t present = lookup[i]; lookup[i] = present ? n : -1; n += present; }
    std::vector<uint64_t> keys((size_t)(m64 ? m64 : 1));
    std::vector<int64_t> off((size_t)T + 1, 0);
    for (int t = 0; t < T; t++) off[t + 1] = off[t] + (int64_t)parts[t].src.size();
#pragma omp parallel for num_threads(T) schedule(static, 1)
    for (int t = 0; t < T; t++) {
        const auto& p = parts[t];
        uint64_t* k = keys.data() + off[t];
        for (size_t i = 0; i < p.src.size(); i++) k[i] = ((uint64_t)(uint32_t)lookup[p.src[i]] << 32) | (uint32_t)lookup[p.dst[i]];
    }
    return gmc_graph_from_edge_keys(out, n, (int32_t)m64, keys.data(), make_reverse);
}
