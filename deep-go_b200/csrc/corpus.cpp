// Packed corpus + native minibatch loader.  Host only, no CUDA.
//
// The reference keeps ONE FILE PER POSITION (3,598-byte Torch7 records) and loads a minibatch by `torch.load`-ing n random
// files in a pool of 32 Lua threads (dataloader.lua:113-125, data.lua:14-19,29-37,82-92).  At the rates the GPU path
// trains at (hundreds of thousands of positions per second) that is 3-4 orders of magnitude short.  Here the same records
// live in one file, fixed 3,264 bytes each, and a minibatch is n memcpys out of a read-only mapping -- the output is exactly
// what dg_model_set_batch_raw takes.
//
//   header (64 B): "DGCORP01", u64 count, u32 record bytes (3264), zero padding
//   record:        u8 raw[9][19][19] (flatten_data planes, dataloader.lua:30-39), x, y, player, rank1, rank2, 10 B padding
#include "../../include/deepgo.h"
#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

void dg_set_error(const char *fmt, ...);

namespace {
const char MAGIC[8] = {'D', 'G', 'C', 'O', 'R', 'P', '0', '1'};
const uint32_t RAW_BYTES = 9 * 19 * 19, REC_BYTES = 3264, HDR_BYTES = 64;

// run f(i) for i in [0, n) on up to nthreads host threads; returns the first error text (empty = ok)
template <typename F> std::string parallel_for(long long n, int nthreads, F f) {
    if (nthreads < 1) nthreads = 1;
    if ((long long)nthreads > n) nthreads = (int)(n > 0 ? n : 1);
    std::atomic<long long> next(0);
    std::atomic<bool> failed(false);
    std::vector<std::string> errs(nthreads);
    auto work = [&](int t) {
        for (;;) {
            const long long lo = next.fetch_add(64);
            if (lo >= n || failed.load()) return;
            const long long hi = lo + 64 < n ? lo + 64 : n;
            for (long long i = lo; i < hi; ++i) {
                std::string e = f(i);
                if (!e.empty()) { errs[t] = e; failed.store(true); return; }
            }
        }
    };
    if (nthreads == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
        for (auto &x : th) x.join();
    }
    for (auto &e : errs) if (!e.empty()) return e;
    return std::string();
}
}  // namespace

struct dg_corpus {
    int fd = -1;
    const uint8_t *map = nullptr;
    size_t bytes = 0;
    long long count = 0;
};

extern "C" int dg_corpus_pack(const char *const *paths, long long n, const char *out_path, int nthreads) {
    if (!paths || !out_path || n < 0) { dg_set_error("dg_corpus_pack: bad argument"); return 1; }
    std::vector<uint8_t> recs((size_t)n * REC_BYTES, 0);
    const std::string err = parallel_for(n, nthreads, [&](long long i) -> std::string {
        uint8_t *r = recs.data() + (size_t)i * REC_BYTES;
        int x, y, player, ranks[2];
        if (dg_t7_read_record(paths[i], r, &x, &y, &player, ranks) != 0) return std::string(dg_last_error());
        if (x < 1 || x > 19 || y < 1 || y > 19 || player < 1 || player > 2 || ranks[0] < 0 || ranks[0] > 255 || ranks[1] < 0 || ranks[1] > 255)
            return std::string("dg_corpus_pack: ") + paths[i] + ": move or ranks out of range";
        r[RAW_BYTES] = (uint8_t)x; r[RAW_BYTES + 1] = (uint8_t)y; r[RAW_BYTES + 2] = (uint8_t)player;
        r[RAW_BYTES + 3] = (uint8_t)ranks[0]; r[RAW_BYTES + 4] = (uint8_t)ranks[1];
        return std::string();
    });
    if (!err.empty()) { dg_set_error("%s", err.c_str()); return 5; }
    FILE *f = fopen(out_path, "wb");
    if (!f) { dg_set_error("dg_corpus_pack: cannot create %s", out_path); return 4; }
    uint8_t hdr[HDR_BYTES] = {0};
    memcpy(hdr, MAGIC, 8);
    const uint64_t cnt = (uint64_t)n;
    memcpy(hdr + 8, &cnt, 8);
    memcpy(hdr + 16, &REC_BYTES, 4);
    const bool ok = fwrite(hdr, 1, HDR_BYTES, f) == HDR_BYTES && fwrite(recs.data(), 1, recs.size(), f) == recs.size();
    if (fclose(f) != 0 || !ok) { dg_set_error("dg_corpus_pack: short write to %s", out_path); return 4; }
    return 0;
}

extern "C" int dg_corpus_open(const char *path, dg_corpus **out) {
    if (!path || !out) { dg_set_error("dg_corpus_open: NULL argument"); return 1; }
    *out = nullptr;
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { dg_set_error("dg_corpus_open: cannot open %s", path); return 4; }
    struct stat st;
    if (fstat(fd, &st) != 0 || (size_t)st.st_size < HDR_BYTES) { close(fd); dg_set_error("dg_corpus_open: %s is too short", path); return 5; }
    void *m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_SHARED, fd, 0);
    if (m == MAP_FAILED) { close(fd); dg_set_error("dg_corpus_open: mmap of %s failed", path); return 4; }
    const uint8_t *p = (const uint8_t *)m;
    uint64_t cnt; uint32_t rb;
    memcpy(&cnt, p + 8, 8); memcpy(&rb, p + 16, 4);
    if (memcmp(p, MAGIC, 8) != 0 || rb != REC_BYTES || (uint64_t)st.st_size != HDR_BYTES + cnt * REC_BYTES) {
        munmap(m, (size_t)st.st_size); close(fd);
        dg_set_error("dg_corpus_open: %s is not a packed corpus (magic, record size or length mismatch)", path);
        return 5;
    }
    dg_corpus *c = new dg_corpus();
    c->fd = fd; c->map = p; c->bytes = (size_t)st.st_size; c->count = (long long)cnt;
    *out = c;
    return 0;
}

extern "C" int dg_corpus_close(dg_corpus *c) {
    if (!c) return 0;
    if (c->map) munmap((void *)c->map, c->bytes);
    if (c->fd >= 0) close(c->fd);
    delete c;
    return 0;
}

extern "C" int dg_corpus_count(dg_corpus *c, long long *out) {
    if (!c || !out) { dg_set_error("dg_corpus_count: NULL argument"); return 1; }
    *out = c->count;
    return 0;
}

extern "C" int dg_corpus_read_batch(dg_corpus *c, const long long *idx, int n, uint8_t *raw, uint8_t *player, uint8_t *rank,
                                    uint8_t *x, uint8_t *y, int nthreads) {
    if (!c || !idx || n < 0 || !raw || !player || !rank) { dg_set_error("dg_corpus_read_batch: bad argument"); return 1; }
    if ((x == nullptr) != (y == nullptr)) { dg_set_error("dg_corpus_read_batch: x and y must both be given or both NULL"); return 1; }
    for (int i = 0; i < n; ++i)
        if (idx[i] < 0 || idx[i] >= c->count) { dg_set_error("dg_corpus_read_batch: record %lld outside 0..%lld", idx[i], c->count - 1); return 1; }
    // a copy of 3.2 KB per record: one core moves ~1.6 M records/s, starting a thread costs as much as ~300 of them
    if (nthreads > n / 2048) nthreads = n / 2048;
    parallel_for(n, nthreads, [&](long long i) -> std::string {
        const uint8_t *r = c->map + HDR_BYTES + (size_t)idx[i] * REC_BYTES;
        memcpy(raw + (size_t)i * RAW_BYTES, r, RAW_BYTES);
        const uint8_t pl = r[RAW_BYTES + 2];
        player[i] = pl;
        rank[i] = r[RAW_BYTES + 3 + (pl - 1)];               // data.ranks[data.move.player] (dataloader.lua:87)
        if (x) { x[i] = r[RAW_BYTES]; y[i] = r[RAW_BYTES + 1]; }
        return std::string();
    });
    return 0;
}
