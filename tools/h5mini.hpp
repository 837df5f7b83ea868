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
#ifndef DLCO_H5MINI_HPP
#define DLCO_H5MINI_HPP
#include <cstdint>
#include <cstdio>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace h5mini {

enum DType { U8 = 0, I32 = 1, F32 = 2, F64 = 3 };
inline size_t dtype_size(DType t) { return t == U8 ? 1 : t == F64 ? 8 : 4; }

struct Error : std::runtime_error {
    explicit Error(const std::string &m) : std::runtime_error(m) {}
};

struct DatasetInfo {
    std::vector<uint64_t> dims, maxdims;
    DType dtype = F32;
    int layout = 0;                   // 1 contiguous, 2 chunked, 0 compact
    uint64_t addr = 0, size = 0;      // contiguous: data address / bytes; chunked: B-tree root
    std::vector<uint32_t> chunk;      // chunk dims (without the trailing element size)
    bool deflate = false;
    uint64_t header = 0;              // object header address
    uint64_t rows() const { return dims.empty() ? 0 : dims[0]; }
    uint64_t row_elems() const { uint64_t n = 1; for (size_t i = 1; i < dims.size(); i++) n *= dims[i]; return n; }
};

// ---- reader: the whole file is mmap-ed; rows are decoded on demand ------------------------
class Reader {
public:
    explicit Reader(const std::string &path);
    ~Reader();
    Reader(const Reader &) = delete;
    Reader &operator=(const Reader &) = delete;
    bool exists(const std::string &name) const { return sets_.count(name) != 0; }
    std::vector<std::string> names() const;
    const DatasetInfo &info(const std::string &name) const;
    // rows [lo, hi) of the leading dimension, all trailing dimensions, into dst
    // (dtype of the file; (hi-lo) * row_elems elements).  nthreads > 1 inflates chunks in parallel.
    void read_rows(const std::string &name, uint64_t lo, uint64_t hi, void *dst, int nthreads = 1) const;
    template <class T> std::vector<T> read_all(const std::string &name, int nthreads = 1) const {
        const DatasetInfo &d = info(name);
        if (dtype_size(d.dtype) != sizeof(T)) throw Error("h5mini: element size mismatch reading " + name);
        std::vector<T> v(d.rows() * d.row_elems());
        if (!v.empty()) read_rows(name, 0, d.rows(), v.data(), nthreads);
        return v;
    }

private:
    struct Chunk { uint64_t addr; uint32_t bytes, mask; uint64_t off[4]; };
    void parse_header(uint64_t addr, DatasetInfo &d) const;
    void walk_group(uint64_t btree, uint64_t heap);
    void collect_chunks(const DatasetInfo &d, uint64_t node, uint64_t lo, uint64_t hi, std::vector<Chunk> &out) const;
    const uint8_t *at(uint64_t off, uint64_t len) const;
    std::string path_;
    const uint8_t *base_ = nullptr;
    uint64_t size_ = 0;
    int fd_ = -1;
    std::map<std::string, DatasetInfo> sets_;
};

// ---- writer: data is streamed, metadata (headers, B-trees, heap, group) is laid down at close
class Writer {
public:
    explicit Writer(const std::string &path);
    ~Writer();
    Writer(const Writer &) = delete;
    Writer &operator=(const Writer &) = delete;
    // whole dataset at once, contiguous layout (pj-learn `W`, `A`)
    void write_contiguous(const std::string &name, DType t, const std::vector<uint64_t> &dims, const void *data);
    // chunked dataset grown along dim 0: chunk = chunk_rows x trailing dims; deflate level 0 = no filter;
    // unlimited = max dim 0 is H5S_UNLIMITED (pr-learn `w`).  Returns a handle for append_rows().
    int create_chunked(const std::string &name, DType t, const std::vector<uint64_t> &trailing_dims,
                       uint32_t chunk_rows, int deflate_level, bool unlimited);
    // append nrows rows; rows are buffered to whole chunks, the last partial chunk is padded with
    // zeros at close (HDF5 stores edge chunks full-size).  nthreads > 1 deflates chunks in parallel.
    void append_rows(int handle, const void *data, uint64_t nrows, int nthreads = 1);
    void close();

private:
    struct ChunkRec { uint64_t addr; uint32_t bytes; uint64_t row; };
    struct Set {
        std::string name; DType t; std::vector<uint64_t> dims; bool chunked = false, unlimited = false;
        uint32_t chunk_rows = 0; int level = 0; uint64_t addr = 0, bytes = 0;
        std::vector<ChunkRec> chunks; std::vector<uint8_t> pending; uint64_t header = 0, btree = 0;
    };
    void put(const void *p, size_t n);
    void pad_to(uint64_t align);
    void flush_chunks(Set &s, bool final, int nthreads);
    uint64_t write_chunk_btree(Set &s);
    uint64_t write_dataset_header(Set &s);
    FILE *f_ = nullptr;
    uint64_t pos_ = 0;
    std::vector<Set> sets_;
    std::string path_;
};

}  // namespace h5mini
#endif
