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
