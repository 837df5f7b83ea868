// Host-side pieces shared by the drop-in tools (comp-fulldists, comp-uprjdists, pr-learn, pj-learn,
// pr-stats, pj-stats): the reference's progress bar, cv::RNG / randShuffle, the train/validation
// split, the ComputePRStats bookkeeping and SelectPRFilters -- restated in plain C++ (the reference
// gets them from OpenCV) -- plus small RAII helpers over the C ABI in include/dlco_b200.h.
// Every numeric step of the hot path runs in libdlco_b200.so; nothing here computes distances,
// margins or gradients.
#ifndef DLCO_TOOLS_COMMON_HPP
#define DLCO_TOOLS_COMMON_HPP
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../include/dlco_b200.h"
#include "h5mini.hpp"

namespace dlco {

// ---- progress bar: "0...10...20... ... 100 - done." in 41 ticks (src/misc.cpp:45-76) ------------
struct Progress {
    int last = -1;
    void update(double done)
    {
        int tick = (int)(done * 40.0);
        tick = std::max(0, std::min(40, tick));
        if (tick < last && last >= 39) last = -1;       // a new run starts
        if (tick <= last) return;
        while (last < tick) {
            ++last;
            if (last % 4 == 0) printf("%d", last / 4 * 10);
            else printf(".");
        }
        if (tick == 40) printf(" - done.\n");
        else fflush(stdout);
    }
};

inline double now_sec()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

inline int host_threads()
{
    const char *e = getenv("DLCO_IO_THREADS");
    int n = e ? atoi(e) : (int)std::thread::hardware_concurrency();
    return n > 0 ? n : 1;
}
inline int deflate_level()
{
    // the reference writes gzip level 9 (e.g. src/comp-fulldists.cpp:290); the level is a storage
    // detail no reader sees, so the default favours write speed.  DLCO_GZIP=9 reproduces it.
    const char *e = getenv("DLCO_GZIP");
    return e ? std::max(0, std::min(9, atoi(e))) : 1;
}
inline int dist_deflate_level()
{
    // "Distance" (f32 distances): deflate buys 10 % on such data (measured: 10.24 GB -> 9.09 GB at level 1, 9.06 GB at
    // level 9) and costs 15 s per GB and 4 threads at ANY level -- 21 s of a 24 s comp-fulldists run on 500k pairs,
    // against 2.8 s of compute (profiles/r02_tool_comp_fulldists.json).  Default: chunked, no filter.  DLCO_GZIP=n
    // (1..9) restores the reference's filter pipeline.
    const char *e = getenv("DLCO_GZIP");
    return e ? std::max(0, std::min(9, atoi(e))) : 0;
}
inline void announce_storage(const char *tool, int level)
{
    fprintf(stderr, "[%s] Distance chunks: %s (DLCO_GZIP=%d); host threads: %d\n", tool,
            level ? "deflate" : "no filter", level, host_threads());
}

// DLCO_MAX_ITERS caps the iteration count of pr-learn / pj-learn / pj-sweep: the workspace
// scripts hard-code -iters; the cap ends a run early at a whole number of steps, logged lines unchanged.
inline unsigned int max_iters_cap(unsigned int n_iter)
{
    const char *e = getenv("DLCO_MAX_ITERS");
    if (!e || !*e) return n_iter;
    const unsigned long long cap = strtoull(e, nullptr, 10);
    if (cap < n_iter) {
        fprintf(stderr, "DLCO_MAX_ITERS=%llu: stopping after %llu of %u iterations\n", cap, cap, n_iter);
        return (unsigned int)cap;
    }
    return n_iter;
}

// ---- cv::RNG (multiply-with-carry) and cv::randShuffle (src/pr-learn.cpp:241-243) -------------
struct CvRng {
    uint64_t state;
    explicit CvRng(uint64_t seed = 0xffffffffu) : state(seed ? seed : 0xffffffffu) {}
    uint32_t next()
    {
        state = (uint64_t)(uint32_t)state * 4164903690u + (state >> 32);
        return (uint32_t)state;
    }
    // RNG::uniform(int a, int b)
    int uniform(int a, int b) { return a == b ? a : a + (int)(next() % (uint32_t)(b - a)); }
};
inline void rand_shuffle(std::vector<int32_t> &v, CvRng &rng)
{
    const size_t n = v.size();
    for (size_t i = 0; i < n; i++) std::swap(v[i], v[rng.next() % (uint32_t)n]);
}

// IdxPos / IdxNeg shuffled with the process-global RNG (default state), positives first; the first
// 80 % of each list is the training part (src/pr-learn.cpp:230-254 == src/pj-learn.cpp:214-237).
struct Split {
    std::vector<int32_t> idx_pos, idx_neg;
    size_t n_pos_trn = 0, n_neg_trn = 0;
    size_t n_pos_val() const { return idx_pos.size() - n_pos_trn; }
    size_t n_neg_val() const { return idx_neg.size() - n_neg_trn; }
};
inline Split split_train_val(const std::vector<uint8_t> &labels)
{
    Split s;
    for (size_t i = 0; i < labels.size(); i++) {
        if (labels[i] == 1) s.idx_pos.push_back((int32_t)i);
        if (labels[i] == 0) s.idx_neg.push_back((int32_t)i);
    }
    CvRng global;
    rand_shuffle(s.idx_pos, global);
    rand_shuffle(s.idx_neg, global);
    const float n_div = 0.80f;
    s.n_pos_trn = (size_t)(s.idx_pos.size() * n_div);
    s.n_neg_trn = (size_t)(s.idx_neg.size() * n_div);
    return s;
}

// ---- ComputePRStats bookkeeping (src/misc.cpp:182-217): rows of PRParams of the rings with w > 0
// that are not all-zero; nzDim = their number; nPR = nzDim - (rows that have a twin) / 2 ----------
struct PrDims { int n_pr, dim, nz_dim; };
inline PrDims pr_dims(const std::vector<float> &pr_params, const float *w, int n_rings, int n_channels = 8)
{
    std::vector<const float *> rows;
    for (int i = 0; i < n_rings; i++) {
        if (!(w[i] > 0.0f)) continue;
        for (int j = 0; j < 8; j++) {
            const float *r = &pr_params[(size_t)(8 * i + j) * 3];
            if (r[0] != 0.0f || r[1] != 0.0f || r[2] != 0.0f) rows.push_back(r);
        }
    }
    int dup = 0;
    for (size_t i = 0; i < rows.size(); i++) {
        bool twin = false;
        for (size_t j = 0; j < rows.size() && !twin; j++)
            twin = i != j && rows[i][0] == rows[j][0] && rows[i][1] == rows[j][1] && rows[i][2] == rows[j][2];
        dup += twin;
    }
    PrDims d;
    d.nz_dim = (int)rows.size();
    d.n_pr = d.nz_dim - dup / 2;
    d.dim = d.n_pr * n_channels;
    return d;
}

// ---- SelectPRFilters (src/misc.cpp:78-169): filters of the rings with w > 0 that are not all-zero,
// first occurrence of each distinct row, then sorted lexicographically -----------------------------
inline std::vector<float> select_pr_filters(const float *filters, int n_filters, int k, const float *w, int *n_sel)
{
    std::vector<const float *> rows;
    for (int i = 0; i < n_filters / 8; i++) {
        if (!(w[i] > 0.0f)) continue;
        for (int j = 0; j < 8; j++) {
            const float *r = filters + (size_t)(8 * i + j) * k;
            bool nz = false;
            for (int c = 0; c < k && !nz; c++) nz = r[c] != 0.0f;
            if (!nz) continue;
            bool seen = false;
            for (size_t u = 0; u < rows.size() && !seen; u++) seen = memcmp(rows[u], r, (size_t)k * sizeof(float)) == 0 ||
                                                                     std::equal(r, r + k, rows[u]);
            if (!seen) rows.push_back(r);
        }
    }
    std::sort(rows.begin(), rows.end(),
              [k](const float *a, const float *b) { return std::lexicographical_compare(a, a + k, b, b + k); });
    std::vector<float> out((size_t)rows.size() * k);
    for (size_t u = 0; u < rows.size(); u++) memcpy(&out[u * k], rows[u], (size_t)k * sizeof(float));
    *n_sel = (int)rows.size();
    return out;
}

// ---- C ABI helpers ------------------------------------------------------------------------------
struct AbiError : std::runtime_error {
    explicit AbiError(const std::string &m) : std::runtime_error(m) {}
};
inline void check(int rc, const char *what)
{
    if (rc != DLCO_OK) throw AbiError(std::string(what) + ": " + dlco_last_error());
}
struct Ctx {
    dlco_ctx *h = nullptr;
    explicit Ctx(int device = 0) { check(dlco_init(device, &h), "dlco_init"); }
    ~Ctx() { if (h) dlco_shutdown(h); }
    Ctx(const Ctx &) = delete;
    Ctx &operator=(const Ctx &) = delete;
    void sync() { check(dlco_sync(h), "dlco_sync"); }
    // "Found GPU: ..." / "Compute Capability: M.m" block of the reference (src/pr-learn.cpp:276-281)
    void print_gpu()
    {
        char name[256];
        int ma = 0, mi = 0;
        check(dlco_device_name(h, name, sizeof(name), &ma, &mi), "dlco_device_name");
        printf("\nFound GPU: %s\nCompute Capability: %d.%d\n\n", name, ma, mi);
        fflush(stdout);
    }
};
template <class T> struct DevBuf {
    Ctx *ctx = nullptr;
    T *p = nullptr;
    size_t n = 0;
    DevBuf() {}
    DevBuf(Ctx &c, size_t count) { alloc(c, count); }
    ~DevBuf() { release(); }
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    void alloc(Ctx &c, size_t count)
    {
        release();
        ctx = &c;
        n = count;
        void *q = nullptr;
        check(dlco_malloc(c.h, std::max<size_t>(count, 1) * sizeof(T), &q), "dlco_malloc");
        p = (T *)q;
    }
    void release()
    {
        if (p) dlco_free(ctx->h, p);
        p = nullptr;
    }
    void upload(const T *src, size_t count, size_t offset = 0)
    {
        if (count) check(dlco_memcpy_h2d(ctx->h, p + offset, src, count * sizeof(T)), "dlco_memcpy_h2d");
    }
    void download(T *dst, size_t count, size_t offset = 0) const
    {
        if (count) check(dlco_memcpy_d2h(ctx->h, dst, p + offset, count * sizeof(T)), "dlco_memcpy_d2h");
    }
};

// "Label" + "Distance" of a comp-fulldists / comp-uprjdists file -> host labels, device matrix.
// Prints the reference's "Load Labels: N" / "Load Distances: N x F" lines and progress bar when
// `verbose` (src/pr-learn.cpp:197-228); pr-stats prints its own prefix instead.
struct DistFile {
    uint64_t n = 0, f = 0;
    std::vector<uint8_t> labels;
};
inline DistFile load_dist_file(const std::string &path, Ctx &ctx, DevBuf<float> &d_dists, bool verbose)
{
    h5mini::Reader r(path);
    const h5mini::DatasetInfo &di = r.info("Distance");
    if (di.dtype != h5mini::F32 || di.dims.size() != 2) throw h5mini::Error("Distance must be a 2-D f32 dataset: " + path);
    const h5mini::DatasetInfo &li = r.info("Label");
    if (li.dtype != h5mini::U8 || li.rows() != di.rows()) throw h5mini::Error("Label must be u8 with one row per pair: " + path);
    DistFile out;
    out.n = di.dims[0];
    out.f = di.dims[1];
    if (verbose) {
        printf("Load Labels: %llu\nLoad Distances: %llu x %llu\n", (unsigned long long)out.n, (unsigned long long)out.n,
               (unsigned long long)out.f);
        fflush(stdout);
    }
    out.labels.resize(out.n);
    if (out.n) r.read_rows("Label", 0, out.n, out.labels.data(), host_threads());
    d_dists.alloc(ctx, out.n * out.f);
    const uint64_t block = std::max<uint64_t>(128, (uint64_t)(256u << 20) / (out.f * 4) / 128 * 128);
    std::vector<float> buf(std::min(block, std::max<uint64_t>(out.n, 1)) * out.f);
    Progress bar;
    for (uint64_t i = 0; i < out.n; i += block) {
        const uint64_t cnt = std::min(block, out.n - i);
        r.read_rows("Distance", i, i + cnt, buf.data(), host_threads());
        d_dists.upload(buf.data(), cnt * out.f, i * out.f);
        bar.update((double)i / (double)out.n);
    }
    bar.update(1.0);
    return out;
}

}  // namespace dlco
#endif
