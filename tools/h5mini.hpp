// h5mini -- the HDF5 on-disk subset the DLCO tools read and write (no libhdf5 in this image).
//
// Covers what OpenCV's hdf module / libhdf5 1.8 produce for the reference's datasets
// (SURVEY.md appendix C): superblock v0 with 8-byte offsets, a symbol-table root group,
// v1 object headers (with continuation blocks), dataspace v1, fixed-point / IEEE
// little-endian datatypes, data layout v3 contiguous or chunked (v1 B-tree index, any depth),
// optional deflate.  Dataset names and shapes: `Indices`, `Patches`
// (src/conv-impatches.cpp:130-134), `PRFilters`, `PRParams`, `RingParams`
// (src/gen-poolregion.cpp:224-233), `Label`, `Distance` (src/comp-fulldists.cpp:232,290),
// `w` (src/pr-learn.cpp:397-407), `W`, `A` (src/pj-learn.cpp:593-594).
//
// The writer lays files out the way libhdf5 does for these tools (same superblock, group
// B-tree + local heap + one symbol node, 272-byte v1 dataset headers with the same message
// sequence), so that any HDF5 reader opens them; chunk shape and deflate level are storage
// details a reader does not see.  Header-only; link with -lz -lpthread.
#ifndef DLCO_H5MINI_HPP
#define DLCO_H5MINI_HPP
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <map>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

namespace h5mini {

enum DType { U8 = 0, I32 = 1, F32 = 2, F64 = 3 };
inline size_t dtype_size(DType t) { return t == U8 ? 1 : t == F64 ? 8 : 4; }
static const uint64_t UNDEF = ~(uint64_t)0;

struct Error : std::runtime_error {
    explicit Error(const std::string &m) : std::runtime_error(m) {}
};

struct DatasetInfo {
    std::vector<uint64_t> dims, maxdims;
    DType dtype = F32;
    int layout = 0;                   // 1 contiguous, 2 chunked, 0 compact
    uint64_t addr = UNDEF, size = 0;  // contiguous: data address / bytes; chunked: B-tree root
    std::vector<uint32_t> chunk;      // chunk dims (without the trailing element size)
    bool deflate = false;
    uint64_t header = 0;              // object header address
    uint64_t rows() const { return dims.empty() ? 0 : dims[0]; }
    uint64_t row_elems() const {
        uint64_t n = 1;
        for (size_t i = 1; i < dims.size(); i++) n *= dims[i];
        return n;
    }
};

template <class F> inline void parallel_for(size_t n, int nthreads, F fn)
{
    if (nthreads <= 1 || n <= 1) {
        for (size_t i = 0; i < n; i++) fn(i);
        return;
    }
    const size_t nt = std::min<size_t>((size_t)nthreads, n);
    std::vector<std::thread> th;
    std::vector<std::string> errs(nt);
    for (size_t t = 0; t < nt; t++)
        th.emplace_back([&, t]() {
            try {
                for (size_t i = t; i < n; i += nt) fn(i);
            } catch (const std::exception &e) {
                errs[t] = e.what();
            }
        });
    for (auto &x : th) x.join();
    for (auto &e : errs)
        if (!e.empty()) throw Error(e);
}

// ---- reader: the whole file is mmap-ed; rows are decoded on demand ------------------------
class Reader {
public:
    explicit Reader(const std::string &path) : path_(path)
    {
        fd_ = ::open(path.c_str(), O_RDONLY);
        if (fd_ < 0) throw Error("h5mini: cannot open " + path);
        struct stat st;
        if (fstat(fd_, &st) != 0 || st.st_size < 96) { ::close(fd_); throw Error("h5mini: not an HDF5 file: " + path); }
        size_ = (uint64_t)st.st_size;
        void *p = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
        if (p == MAP_FAILED) { ::close(fd_); throw Error("h5mini: mmap failed on " + path); }
        base_ = (const uint8_t *)p;
        static const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
        if (memcmp(base_, sig, 8) != 0) fail("bad signature");
        if (base_[8] != 0 || base_[13] != 8 || base_[14] != 8) fail("unsupported superblock (need v0, 8-byte offsets)");
        const uint64_t ohdr = u64(56 + 8);
        const uint32_t ctype = u32(56 + 16);
        uint64_t btree, heap;
        if (ctype == 1) {
            btree = u64(80);
            heap = u64(88);
        } else {
            bool found = false;
            for (auto &m : messages(ohdr))
                if (m.type == 0x11) { btree = rd64(m.p); heap = rd64(m.p + 8); found = true; }
            if (!found) fail("root group has no symbol table");
        }
        walk_group(btree, heap);
    }
    ~Reader()
    {
        if (base_) munmap((void *)base_, size_);
        if (fd_ >= 0) ::close(fd_);
    }
    Reader(const Reader &) = delete;
    Reader &operator=(const Reader &) = delete;

    bool exists(const std::string &name) const { return sets_.count(name) != 0; }
    std::vector<std::string> names() const
    {
        std::vector<std::string> v;
        for (auto &kv : sets_) v.push_back(kv.first);
        return v;
    }
    const DatasetInfo &info(const std::string &name) const
    {
        auto it = sets_.find(name);
        if (it == sets_.end()) throw Error("h5mini: no dataset '" + name + "' in " + path_);
        return it->second;
    }

    // rows [lo, hi) of the leading dimension, all trailing dimensions, into dst (file dtype).
    void read_rows(const std::string &name, uint64_t lo, uint64_t hi, void *dst, int nthreads = 1) const
    {
        const DatasetInfo &d = info(name);
        if (lo > hi || hi > d.rows()) throw Error("h5mini: row range out of bounds reading " + name);
        const size_t es = dtype_size(d.dtype);
        const uint64_t re = d.row_elems();
        if (hi == lo) return;
        if (d.layout == 1) {
            if (d.addr == UNDEF) { memset(dst, 0, (hi - lo) * re * es); return; }
            memcpy(dst, at(d.addr + lo * re * es, (hi - lo) * re * es), (hi - lo) * re * es);
            return;
        }
        if (d.layout != 2) throw Error("h5mini: unsupported layout of " + name);
        memset(dst, 0, (hi - lo) * re * es);
        if (d.addr == UNDEF) return;
        std::vector<Chunk> chunks;
        collect_chunks(d, d.addr, lo, hi, chunks);
        const size_t rank = d.dims.size();
        uint64_t celems = 1;
        for (auto c : d.chunk) celems *= c;
        parallel_for(chunks.size(), nthreads, [&](size_t ci) {
            const Chunk &c = chunks[ci];
            std::vector<uint8_t> tmp;
            const uint8_t *raw = at(c.addr, c.bytes);
            if (d.deflate && !(c.mask & 1)) {
                tmp.resize(celems * es);
                uLongf n = (uLongf)tmp.size();
                if (uncompress(tmp.data(), &n, raw, c.bytes) != Z_OK || n != tmp.size())
                    throw Error("h5mini: inflate failed in " + name);
                raw = tmp.data();
            } else if (c.bytes < celems * es) {
                throw Error("h5mini: short chunk in " + name);
            }
            copy_chunk(d, c, raw, lo, hi, (uint8_t *)dst, rank, es);
        });
    }
    template <class T> std::vector<T> read_all(const std::string &name, int nthreads = 1) const
    {
        const DatasetInfo &d = info(name);
        if (dtype_size(d.dtype) != sizeof(T)) throw Error("h5mini: element size mismatch reading " + name);
        std::vector<T> v(d.rows() * d.row_elems());
        if (!v.empty()) read_rows(name, 0, d.rows(), v.data(), nthreads);
        return v;
    }

private:
    struct Chunk { uint64_t addr; uint32_t bytes, mask; uint64_t off[5]; };
    struct Msg { uint16_t type; uint16_t size; const uint8_t *p; };

    [[noreturn]] void fail(const std::string &m) const { throw Error("h5mini: " + path_ + ": " + m); }
    const uint8_t *at(uint64_t off, uint64_t len) const
    {
        if (off > size_ || len > size_ - off) fail("address out of file bounds");
        return base_ + off;
    }
    static uint16_t rd16(const uint8_t *p) { uint16_t v; memcpy(&v, p, 2); return v; }
    static uint32_t rd32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
    static uint64_t rd64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; }
    uint32_t u32(uint64_t off) const { return rd32(at(off, 4)); }
    uint64_t u64(uint64_t off) const { return rd64(at(off, 8)); }

    std::vector<Msg> messages(uint64_t addr) const
    {
        const uint8_t *h = at(addr, 16);
        if (h[0] != 1) fail("unsupported object header version");
        const unsigned nmsg = rd16(h + 2);
        std::vector<std::pair<uint64_t, uint64_t>> blocks{{addr + 16, rd32(h + 8)}};
        std::vector<Msg> out;
        for (size_t bi = 0; bi < blocks.size() && out.size() < nmsg; bi++) {
            uint64_t p = blocks[bi].first;
            const uint64_t end = p + blocks[bi].second;
            while (p + 8 <= end && out.size() < nmsg) {
                const uint8_t *m = at(p, 8);
                Msg msg{rd16(m), rd16(m + 2), nullptr};
                msg.p = at(p + 8, msg.size);
                p += 8 + msg.size;
                if (msg.type == 0x10) blocks.push_back({rd64(msg.p), rd64(msg.p + 8)});
                out.push_back(msg);
            }
        }
        return out;
    }

    void parse_header(uint64_t addr, DatasetInfo &d) const
    {
        d.header = addr;
        bool have_space = false, have_type = false, have_layout = false;
        for (auto &m : messages(addr)) {
            const uint8_t *p = m.p;
            if (m.type == 0x01) {
                if (p[0] != 1) fail("unsupported dataspace version");
                const unsigned rank = p[1];
                d.dims.resize(rank);
                for (unsigned i = 0; i < rank; i++) d.dims[i] = rd64(p + 8 + 8 * i);
                d.maxdims = d.dims;
                if (p[2] & 1)
                    for (unsigned i = 0; i < rank; i++) d.maxdims[i] = rd64(p + 8 + 8 * rank + 8 * i);
                have_space = true;
            } else if (m.type == 0x03) {
                const unsigned cls = p[0] & 0x0f;
                const uint32_t size = rd32(p + 4);
                if (cls == 0 && size == 1) d.dtype = U8;
                else if (cls == 0 && size == 4) d.dtype = I32;
                else if (cls == 1 && size == 4) d.dtype = F32;
                else if (cls == 1 && size == 8) d.dtype = F64;
                else fail("unsupported datatype");
                if (p[1] & 1) fail("big-endian data is not supported");
                have_type = true;
            } else if (m.type == 0x08) {
                if (p[0] != 3) fail("unsupported data layout version");
                d.layout = p[1];
                if (d.layout == 1) {
                    d.addr = rd64(p + 2);
                    d.size = rd64(p + 10);
                } else if (d.layout == 2) {
                    const unsigned nd = p[2];
                    d.addr = rd64(p + 3);
                    d.chunk.resize(nd - 1);
                    for (unsigned i = 0; i + 1 < nd; i++) d.chunk[i] = rd32(p + 11 + 4 * i);
                } else {
                    fail("compact layout is not supported");
                }
                have_layout = true;
            } else if (m.type == 0x0b) {
                if (p[0] != 1) fail("unsupported filter pipeline version");
                const unsigned nf = p[1];
                const uint8_t *q = p + 8;
                for (unsigned i = 0; i < nf; i++) {
                    const unsigned id = rd16(q), nlen = rd16(q + 2), ncv = rd16(q + 6);
                    q += 8 + (nlen + 7) / 8 * 8 + 4 * ncv + (ncv % 2 ? 4 : 0);
                    if (id == 1) d.deflate = true;
                    else fail("unsupported filter (only deflate)");
                }
            }
        }
        if (!have_space || !have_type || !have_layout) fail("incomplete dataset header");
        if (d.layout == 2 && d.chunk.size() != d.dims.size()) fail("chunk rank mismatch");
        if (d.dims.size() > 4) fail("rank > 4 is not supported");
    }

    void walk_group(uint64_t btree, uint64_t heap)
    {
        const uint8_t *hp = at(heap, 32);
        if (memcmp(hp, "HEAP", 4) != 0) fail("bad local heap");
        const uint64_t hsize = rd64(hp + 8), hdata = rd64(hp + 24);
        const char *names = (const char *)at(hdata, hsize);
        walk_group_node(btree, names, hsize);
    }
    void walk_group_node(uint64_t addr, const char *names, uint64_t hsize)
    {
        const uint8_t *n = at(addr, 24);
        if (memcmp(n, "TREE", 4) != 0 || n[4] != 0) fail("bad group B-tree node");
        const unsigned level = n[5], used = rd16(n + 6);
        for (unsigned i = 0; i < used; i++) {
            const uint64_t child = u64(addr + 24 + 8 + 16 * (uint64_t)i);
            if (level > 0) { walk_group_node(child, names, hsize); continue; }
            const uint8_t *s = at(child, 8);
            if (memcmp(s, "SNOD", 4) != 0) fail("bad symbol node");
            const unsigned ns = rd16(s + 6);
            for (unsigned k = 0; k < ns; k++) {
                const uint8_t *e = at(child + 8 + 40 * (uint64_t)k, 40);
                const uint64_t noff = rd64(e), ohdr = rd64(e + 8);
                if (noff >= hsize) fail("bad link name offset");
                std::string name(names + noff, strnlen(names + noff, hsize - noff));
                DatasetInfo d;
                try {
                    parse_header(ohdr, d);
                } catch (const Error &) {
                    continue;      // not a dataset we understand (e.g. a sub-group): skip
                }
                sets_[name] = d;
            }
        }
    }

    void collect_chunks(const DatasetInfo &d, uint64_t node, uint64_t lo, uint64_t hi, std::vector<Chunk> &out) const
    {
        const size_t nd = d.dims.size() + 1;
        const uint64_t ksz = 8 + 8 * nd;
        const uint8_t *n = at(node, 24);
        if (memcmp(n, "TREE", 4) != 0 || n[4] != 1) fail("bad chunk B-tree node");
        const unsigned level = n[5], used = rd16(n + 6);
        const uint8_t *p = at(node + 24, (uint64_t)used * (ksz + 8) + ksz);
        for (unsigned i = 0; i < used; i++, p += ksz + 8) {
            const uint64_t row0 = rd64(p + 8), next_row0 = rd64(p + ksz + 8 + 8);
            const uint64_t child = rd64(p + ksz);
            if (level > 0) {
                // subtree i holds keys in [key_i, key_{i+1}]: prune on the leading dimension
                if (row0 >= hi && i > 0) break;
                if (next_row0 + d.chunk[0] <= lo && next_row0 >= row0 && i + 1 < used) continue;
                collect_chunks(d, child, lo, hi, out);
                continue;
            }
            if (row0 >= hi || row0 + d.chunk[0] <= lo) continue;
            Chunk c;
            c.addr = child;
            c.bytes = rd32(p);
            c.mask = rd32(p + 4);
            for (size_t k = 0; k < nd && k < 5; k++) c.off[k] = rd64(p + 8 + 8 * k);
            out.push_back(c);
        }
    }

    static void copy_chunk(const DatasetInfo &d, const Chunk &c, const uint8_t *raw, uint64_t lo, uint64_t hi,
                           uint8_t *dst, size_t rank, size_t es)
    {
        // iterate over the chunk rows of all but the last dimension; copy runs along the last
        uint64_t cd[4] = {1, 1, 1, 1}, dd[4] = {1, 1, 1, 1}, off[4] = {0, 0, 0, 0};
        for (size_t k = 0; k < rank; k++) { cd[4 - rank + k] = d.chunk[k]; dd[4 - rank + k] = d.dims[k]; off[4 - rank + k] = c.off[k]; }
        const size_t lead = 4 - rank;          // index of the dataset's dimension 0
        const uint64_t last_n = off[3] >= dd[3] ? 0 : std::min<uint64_t>(cd[3], dd[3] - off[3]);
        if (!last_n) return;
        for (uint64_t i0 = 0; i0 < cd[0]; i0++)
            for (uint64_t i1 = 0; i1 < cd[1]; i1++)
                for (uint64_t i2 = 0; i2 < cd[2]; i2++) {
                    const uint64_t g[3] = {off[0] + i0, off[1] + i1, off[2] + i2};
                    if (g[0] >= dd[0] || g[1] >= dd[1] || g[2] >= dd[2]) continue;
                    uint64_t gl[4] = {g[0], g[1], g[2], off[3]};
                    // leading-dimension window [lo, hi)
                    uint64_t first = 0, n = last_n;
                    if (lead == 3) {            // rank 1: the run itself is along dimension 0
                        const uint64_t a = std::max<uint64_t>(off[3], lo), b = std::min<uint64_t>(off[3] + last_n, hi);
                        if (a >= b) continue;
                        first = a - off[3];
                        n = b - a;
                        gl[3] = a - lo;
                    } else {
                        if (gl[lead] < lo || gl[lead] >= hi) continue;
                        gl[lead] -= lo;
                    }
                    uint64_t dims_w[4] = {dd[0], dd[1], dd[2], dd[3]};
                    dims_w[lead] = hi - lo;
                    const uint64_t o = ((gl[0] * dims_w[1] + gl[1]) * dims_w[2] + gl[2]) * dims_w[3] + gl[3];
                    const uint64_t s = ((i0 * cd[1] + i1) * cd[2] + i2) * cd[3] + first;
                    memcpy(dst + o * es, raw + s * es, n * es);
                }
    }

    std::string path_;
    const uint8_t *base_ = nullptr;
    uint64_t size_ = 0;
    int fd_ = -1;
    std::map<std::string, DatasetInfo> sets_;
};

// ---- writer ----------------------------------------------------------------------------------
// File layout: [0, META) metadata region (superblock, root group header, group B-tree, local heap,
// dataset headers, symbol node), rewritten by flush(); data and chunk B-tree nodes are appended
// behind it.  flush() leaves a complete, readable file, so an appendable dataset (`w`) can be
// flushed after every row like the reference's dsinsert (src/pr-learn.cpp:397-407).
class Writer {
public:
    explicit Writer(const std::string &path) : path_(path)
    {
        f_ = fopen(path.c_str(), "wb+");
        if (!f_) throw Error("h5mini: cannot create " + path);
        std::vector<uint8_t> z(META, 0);
        if (fwrite(z.data(), 1, META, f_) != META) throw Error("h5mini: write failed: " + path);
        pos_ = META;
        mtime_ = (uint32_t)time(nullptr);
    }
    ~Writer()
    {
        try { close(); } catch (...) {}
    }
    Writer(const Writer &) = delete;
    Writer &operator=(const Writer &) = delete;

    // whole dataset at once, contiguous layout (pj-learn `W`, `A`: src/pj-learn.cpp:593-594)
    void write_contiguous(const std::string &name, DType t, const std::vector<uint64_t> &dims, const void *data)
    {
        Set s;
        s.name = name; s.t = t; s.dims = dims; s.maxdims = dims; s.chunked = false;
        uint64_t n = dtype_size(t);
        for (auto d : dims) n *= d;
        s.bytes = n;
        pad_to(8);
        s.addr = n ? pos_ : UNDEF;
        put(data, n);
        add_set(s);
    }
    // chunked dataset grown along dimension 0: chunk = chunk_rows x trailing dims; deflate_level 0 = no
    // filter; unlimited: max dim 0 = H5S_UNLIMITED (pr-learn `w`).  Returns a handle for append_rows().
    int create_chunked(const std::string &name, DType t, const std::vector<uint64_t> &trailing_dims,
                       uint32_t chunk_rows, int deflate_level, bool unlimited)
    {
        if (!chunk_rows) throw Error("h5mini: chunk_rows must be positive");
        Set s;
        s.name = name; s.t = t; s.chunked = true; s.unlimited = unlimited;
        s.chunk_rows = chunk_rows; s.level = deflate_level;
        s.dims.push_back(0);
        for (auto d : trailing_dims) s.dims.push_back(d);
        s.maxdims = s.dims;
        s.row_bytes = dtype_size(t);
        for (auto d : trailing_dims) s.row_bytes *= d;
        add_set(s);
        return (int)sets_.size() - 1;
    }
    // append nrows rows; whole chunks are written at once (nthreads > 1 deflates them in parallel), a
    // trailing partial chunk stays buffered until more rows arrive or the file is flushed / closed
    // (HDF5 stores edge chunks full-size: the padding is zero).
    void append_rows(int handle, const void *data, uint64_t nrows, int nthreads = 1)
    {
        Set &s = sets_.at((size_t)handle);
        if (!s.chunked) throw Error("h5mini: append to a contiguous dataset");
        const uint8_t *src = (const uint8_t *)data;
        const uint64_t cb = (uint64_t)s.chunk_rows * s.row_bytes;
        uint64_t left = nrows * s.row_bytes;
        if (s.flushed_partial) {            // the buffered edge chunk was already written by flush(): redo it
            s.chunks.pop_back();
            s.flushed_partial = false;
        }
        if (!s.pending.empty()) {
            const uint64_t take = std::min<uint64_t>(left, cb - s.pending.size());
            const size_t have = s.pending.size();
            s.pending.resize(have + take);
            memcpy(s.pending.data() + have, src, take);
            src += take; left -= take;
            if (s.pending.size() == cb) {
                write_chunks(s, s.pending.data(), 1, nthreads);
                s.pending.clear();
            }
        }
        const uint64_t whole = left / cb;
        if (whole) {
            write_chunks(s, src, whole, nthreads);
            src += whole * cb; left -= whole * cb;
        }
        if (left) s.pending.assign(src, src + left);
        s.dims[0] += nrows;
        if (!s.unlimited) s.maxdims[0] = s.dims[0];
    }
    // make the file on disk complete and consistent (metadata + chunk indices)
    void flush()
    {
        if (!f_) return;
        for (auto &s : sets_) {
            if (!s.chunked) continue;
            if (!s.pending.empty() && !s.flushed_partial) {
                std::vector<uint8_t> full((size_t)s.chunk_rows * s.row_bytes, 0);
                memcpy(full.data(), s.pending.data(), s.pending.size());
                write_chunks(s, full.data(), 1, 1);
                s.flushed_partial = true;
            }
            s.btree = s.chunks.empty() ? UNDEF : write_chunk_btree(s);
        }
        write_metadata();
        fflush(f_);
    }
    void close()
    {
        if (!f_) return;
        flush();
        fclose(f_);
        f_ = nullptr;
    }

private:
    static const size_t META = 4096;
    static const size_t MAX_SETS = 8;         // one symbol node (2 * leaf K = 8 entries)
    struct ChunkRec { uint64_t addr; uint32_t bytes; uint64_t row; };
    struct Set {
        std::string name; DType t = F32; std::vector<uint64_t> dims, maxdims;
        bool chunked = false, unlimited = false, flushed_partial = false;
        uint32_t chunk_rows = 0; int level = 0; uint64_t row_bytes = 0;
        uint64_t addr = UNDEF, bytes = 0, btree = UNDEF;
        std::vector<ChunkRec> chunks; std::vector<uint8_t> pending;
    };
    void add_set(const Set &s)
    {
        if (sets_.size() >= MAX_SETS) throw Error("h5mini: too many datasets in one file");
        for (auto &o : sets_)
            if (o.name == s.name) throw Error("h5mini: duplicate dataset " + s.name);
        sets_.push_back(s);
    }
    void put(const void *p, size_t n)
    {
        if (n && fwrite(p, 1, n, f_) != n) throw Error("h5mini: write failed: " + path_);
        pos_ += n;
    }
    void pad_to(uint64_t align)
    {
        static const uint8_t z[8] = {0};
        while (pos_ % align) put(z, 1);
    }
    void write_chunks(Set &s, const uint8_t *src, uint64_t n, int nthreads)
    {
        const uint64_t cb = (uint64_t)s.chunk_rows * s.row_bytes;
        const uint64_t row0 = s.chunks.empty() ? 0 : s.chunks.back().row + s.chunk_rows;
        if (s.level <= 0) {
            for (uint64_t i = 0; i < n; i++) {
                s.chunks.push_back({pos_, (uint32_t)cb, row0 + i * s.chunk_rows});
                put(src + i * cb, cb);
            }
            return;
        }
        const uint64_t group = (uint64_t)std::max(nthreads, 1) * 4;
        for (uint64_t g0 = 0; g0 < n; g0 += group) {
            const uint64_t g1 = std::min(n, g0 + group);
            std::vector<std::vector<uint8_t>> z(g1 - g0);
            parallel_for((size_t)(g1 - g0), nthreads, [&](size_t i) {
                uLongf len = compressBound((uLong)cb);
                z[i].resize(len);
                if (compress2(z[i].data(), &len, src + (g0 + i) * cb, (uLong)cb, s.level) != Z_OK)
                    throw Error("h5mini: deflate failed");
                z[i].resize(len);
            });
            for (uint64_t i = g0; i < g1; i++) {
                s.chunks.push_back({pos_, (uint32_t)z[i - g0].size(), row0 + i * s.chunk_rows});
                put(z[i - g0].data(), z[i - g0].size());
            }
        }
    }
    static void w16(std::vector<uint8_t> &b, size_t o, uint16_t v) { memcpy(&b[o], &v, 2); }
    static void w32(std::vector<uint8_t> &b, size_t o, uint32_t v) { memcpy(&b[o], &v, 4); }
    static void w64(std::vector<uint8_t> &b, size_t o, uint64_t v) { memcpy(&b[o], &v, 8); }

    // v1 B-tree (node type 1) over the chunks of s, 64 entries per node; returns the root address
    uint64_t write_chunk_btree(Set &s)
    {
        const size_t nd = s.dims.size() + 1, ksz = 8 + 8 * nd, K2 = 64;
        const size_t node_bytes = 24 + (K2 + 1) * ksz + K2 * 8;
        struct Ent { uint64_t child; uint32_t bytes; uint64_t row; };
        std::vector<Ent> level;
        for (auto &c : s.chunks) level.push_back({c.addr, c.bytes, c.row});
        const uint64_t end_row = s.chunks.back().row + s.chunk_rows;
        unsigned depth = 0;
        pad_to(8);
        for (;;) {
            const size_t nn = (level.size() + K2 - 1) / K2;
            const uint64_t base = pos_;
            std::vector<Ent> up;
            for (size_t n = 0; n < nn; n++) {
                const size_t a = n * K2, b = std::min(level.size(), a + K2);
                std::vector<uint8_t> nb(node_bytes, 0);
                memcpy(&nb[0], "TREE", 4);
                nb[4] = 1;
                nb[5] = (uint8_t)depth;
                w16(nb, 6, (uint16_t)(b - a));
                w64(nb, 8, n ? base + (n - 1) * node_bytes : UNDEF);
                w64(nb, 16, n + 1 < nn ? base + (n + 1) * node_bytes : UNDEF);
                size_t o = 24;
                for (size_t i = a; i < b; i++) {
                    w32(nb, o, level[i].bytes);          // key i: size, filter mask 0, offsets
                    w64(nb, o + 8, level[i].row);
                    w64(nb, o + ksz, level[i].child);
                    o += ksz + 8;
                }
                // final key: where the next subtree starts (one chunk past the end for the last node)
                const uint64_t next_row = b < level.size() ? level[b].row : end_row;
                w64(nb, o + 8, next_row);
                if (b == level.size()) {          // libhdf5's right-most key: one chunk past the end in EVERY dimension,
                    for (size_t k = 1; k + 1 < nd; k++) w64(nb, o + 8 + 8 * k, s.dims[k]);      // element size included
                    w64(nb, o + 8 + 8 * (nd - 1), dtype_size(s.t));
                }
                put(nb.data(), nb.size());
                up.push_back({base + n * node_bytes, level[a].bytes, level[a].row});
            }
            if (nn == 1) return base;
            level.swap(up);
            depth++;
        }
    }

    static std::vector<uint8_t> msg(uint16_t type, const std::vector<uint8_t> &body, uint8_t flags = 0)
    {
        std::vector<uint8_t> m(8 + (body.size() + 7) / 8 * 8, 0);
        w16(m, 0, type);
        w16(m, 2, (uint16_t)(m.size() - 8));
        m[4] = flags;
        memcpy(&m[8], body.data(), body.size());
        return m;
    }
    std::vector<uint8_t> dataset_header(const Set &s) const
    {
        const size_t rank = s.dims.size();
        std::vector<std::vector<uint8_t>> msgs;
        {   // dataspace v1 with max dims
            std::vector<uint8_t> b(8 + 16 * rank, 0);
            b[0] = 1; b[1] = (uint8_t)rank; b[2] = 1;
            for (size_t i = 0; i < rank; i++) {
                w64(b, 8 + 8 * i, s.dims[i]);
                w64(b, 8 + 8 * rank + 8 * i, (i == 0 && s.unlimited) ? UNDEF : s.maxdims[i]);
            }
            msgs.push_back(msg(0x01, b));
        }
        {   // datatype v1
            std::vector<uint8_t> b;
            if (s.t == F32) b = {0x11, 0x20, 0x1f, 0x00, 4, 0, 0, 0, 0, 0, 32, 0, 23, 8, 0, 23, 127, 0, 0, 0};
            else if (s.t == F64) b = {0x11, 0x20, 0x3f, 0x00, 8, 0, 0, 0, 0, 0, 64, 0, 52, 11, 0, 52, 0xff, 0x03, 0, 0};
            else if (s.t == I32) b = {0x10, 0x08, 0x00, 0x00, 4, 0, 0, 0, 0, 0, 32, 0};
            else b = {0x10, 0x00, 0x00, 0x00, 1, 0, 0, 0, 0, 0, 8, 0};
            msgs.push_back(msg(0x03, b, 1));
        }
        // fill value v2: allocation time (3 incremental for chunked, 2 late), write time if-set, default value
        msgs.push_back(msg(0x05, {2, (uint8_t)(s.chunked ? 3 : 2), 2, 1, 0, 0, 0, 0}));
        if (s.chunked && s.level > 0) {
            std::vector<uint8_t> b = {1, 1, 0, 0, 0, 0, 0, 0, 1, 0, 8, 0, 1, 0, 1, 0, 'd', 'e', 'f', 'l', 'a', 't', 'e', 0,
                                      (uint8_t)s.level, 0, 0, 0, 0, 0, 0, 0};
            msgs.push_back(msg(0x0b, b));
        }
        if (s.chunked) {
            std::vector<uint8_t> b(11 + 4 * (rank + 1), 0);
            b[0] = 3; b[1] = 2; b[2] = (uint8_t)(rank + 1);
            w64(b, 3, s.btree);
            w32(b, 11, s.chunk_rows);
            for (size_t i = 1; i < rank; i++) w32(b, 11 + 4 * i, (uint32_t)s.dims[i]);
            w32(b, 11 + 4 * rank, (uint32_t)dtype_size(s.t));
            msgs.push_back(msg(0x08, b));
        } else {
            std::vector<uint8_t> b(18, 0);
            b[0] = 3; b[1] = 1;
            w64(b, 2, s.addr);
            w64(b, 10, s.bytes);
            msgs.push_back(msg(0x08, b));
        }
        {
            std::vector<uint8_t> b(8, 0);
            b[0] = 1;
            w32(b, 4, mtime_);
            msgs.push_back(msg(0x12, b));
        }
        size_t used = 0;
        for (auto &m : msgs) used += m.size();
        const size_t HDR = 256;               // libhdf5's default dataset header allocation
        if (used + 8 > HDR) throw Error("h5mini: dataset header overflow");
        msgs.push_back(msg(0x00, std::vector<uint8_t>(HDR - used - 8, 0)));
        std::vector<uint8_t> h(16, 0);
        h[0] = 1;
        w16(h, 2, (uint16_t)msgs.size());
        w32(h, 4, 1);
        w32(h, 8, (uint32_t)HDR);
        for (auto &m : msgs) h.insert(h.end(), m.begin(), m.end());
        return h;
    }

    void write_metadata()
    {
        std::vector<uint8_t> b(META, 0);
        // local heap data: "" at 0, then the names (8-byte aligned), then one free block
        std::vector<size_t> order(sets_.size());
        for (size_t i = 0; i < order.size(); i++) order[i] = i;
        std::sort(order.begin(), order.end(), [&](size_t x, size_t y) { return sets_[x].name < sets_[y].name; });
        std::vector<uint64_t> name_off(sets_.size());
        size_t hp = 8;
        for (size_t i = 0; i < sets_.size(); i++) {
            name_off[i] = hp;
            hp += (sets_[i].name.size() + 1 + 7) / 8 * 8;
        }
        const size_t heap_data = std::max<size_t>(88, hp + 16);
        const size_t ROOT = 96, BTREE = 136, HEAP = 680, HDATA = 712;
        const size_t first_hdr = HDATA + (heap_data + 7) / 8 * 8;
        const size_t snod = first_hdr + 272 * sets_.size();
        if (snod + 328 > META) throw Error("h5mini: metadata region overflow");
        // superblock v0
        static const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
        memcpy(&b[0], sig, 8);
        b[13] = 8; b[14] = 8;
        w16(b, 16, 4); w16(b, 18, 16);
        w64(b, 24, 0); w64(b, 32, UNDEF); w64(b, 40, pos_); w64(b, 48, UNDEF);
        w64(b, 56, 0); w64(b, 64, ROOT); w32(b, 72, 1); w64(b, 80, BTREE); w64(b, 88, HEAP);
        // root group object header: one symbol-table message
        b[ROOT] = 1; w16(b, ROOT + 2, 1); w32(b, ROOT + 4, 1); w32(b, ROOT + 8, 24);
        w16(b, ROOT + 16, 0x11); w16(b, ROOT + 18, 16);
        w64(b, ROOT + 24, BTREE); w64(b, ROOT + 32, HEAP);
        // group B-tree: one leaf entry -> the symbol node
        memcpy(&b[BTREE], "TREE", 4);
        w16(b, BTREE + 6, sets_.empty() ? 0 : 1);
        w64(b, BTREE + 8, UNDEF); w64(b, BTREE + 16, UNDEF);
        w64(b, BTREE + 24, 0);
        w64(b, BTREE + 32, snod);
        w64(b, BTREE + 40, sets_.empty() ? 0 : name_off[order.back()]);
        // local heap
        memcpy(&b[HEAP], "HEAP", 4);
        w64(b, HEAP + 8, heap_data); w64(b, HEAP + 16, hp); w64(b, HEAP + 24, HDATA);
        for (size_t i = 0; i < sets_.size(); i++) memcpy(&b[HDATA + name_off[i]], sets_[i].name.c_str(), sets_[i].name.size());
        w64(b, HDATA + hp, 1);                               // H5HL_FREE_NULL
        w64(b, HDATA + hp + 8, heap_data - hp);
        // dataset headers + symbol node (entries sorted by name)
        for (size_t i = 0; i < sets_.size(); i++) {
            std::vector<uint8_t> h = dataset_header(sets_[i]);
            memcpy(&b[first_hdr + 272 * i], h.data(), h.size());
        }
        memcpy(&b[snod], "SNOD", 4);
        b[snod + 4] = 1;
        w16(b, snod + 6, (uint16_t)sets_.size());
        for (size_t k = 0; k < order.size(); k++) {
            const size_t e = snod + 8 + 40 * k;
            w64(b, e, name_off[order[k]]);
            w64(b, e + 8, first_hdr + 272 * order[k]);
        }
        if (fseek(f_, 0, SEEK_SET) != 0 || fwrite(b.data(), 1, META, f_) != META || fseek(f_, (long)pos_, SEEK_SET) != 0)
            throw Error("h5mini: metadata write failed: " + path_);
    }

    FILE *f_ = nullptr;
    uint64_t pos_ = 0;
    uint32_t mtime_ = 0;
    std::vector<Set> sets_;
    std::string path_;
};

}  // namespace h5mini
#endif
