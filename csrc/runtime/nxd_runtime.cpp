// Host-side native runtime of the inference path (no GPU code, plain C++17, built with g++):
//
//   * safetensors reader: parses the header, mmaps the file and copies any set of tensors into caller-provided
//     buffers with a pool of threads (the reference loads `tp{r}_sharded_checkpoint.safetensors` through the Python
//     safetensors package, one tensor at a time: examples/inference/runner.py + trace/model_builder.py:302-321).  The
//     destination is usually pinned host memory, from where one cudaMemcpyAsync per tensor moves it to the GPU.
//     A row-range variant copies only rows [r0, r1) of a 2-D tensor, which is what slicing a full checkpoint for a
//     tensor-parallel rank needs for dim-0 sharded weights.
//   * bucket router: the (model key, batch, sequence) -> compiled-graph lookup of NxDModel
//     (trace/model_builder.py:64-120 `NxDModelExecutor` + `router`): smallest bucket that covers the request.
//
// C ABI, thread-safe per handle, errors through nxd_rt_last_error().
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define NXD_RT_API extern "C" __attribute__((visibility("default")))

namespace {

thread_local char g_err[512] = "";

int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return -1;
}

struct TensorInfo {
    std::string name, dtype;
    std::vector<int64_t> shape;
    uint64_t begin = 0, end = 0;   // byte offsets inside the data section
};

struct StFile {
    int fd = -1;
    uint8_t* map = nullptr;
    size_t size = 0;
    size_t data_off = 0;
    std::vector<TensorInfo> tensors;
    std::map<std::string, int> index;
};

// ---- minimal JSON scanner for the safetensors header (objects of strings / integer arrays / nested objects) ----
struct Scanner {
    const char* p;
    const char* e;
    void ws() { while (p < e && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
    bool eat(char c) { ws(); if (p < e && *p == c) { ++p; return true; } return false; }
    bool str(std::string& out) {
        ws();
        if (p >= e || *p != '"') return false;
        ++p;
        out.clear();
        while (p < e && *p != '"') {
            if (*p == '\\' && p + 1 < e) {
                ++p;
                switch (*p) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'u': out.push_back('?'); p += 4; break;   // names are ASCII in practice
                    default: out.push_back(*p);
                }
                ++p;
            } else {
                out.push_back(*p++);
            }
        }
        if (p >= e) return false;
        ++p;
        return true;
    }
    bool integer(int64_t& v) {
        ws();
        const char* s = p;
        if (p < e && *p == '-') ++p;
        while (p < e && *p >= '0' && *p <= '9') ++p;
        if (p == s) return false;
        v = std::strtoll(std::string(s, p).c_str(), nullptr, 10);
        return true;
    }
    bool skip_value() {   // anything: used for __metadata__
        ws();
        if (p >= e) return false;
        if (*p == '"') { std::string t; return str(t); }
        if (*p == '{' || *p == '[') {
            const char open = *p, close = open == '{' ? '}' : ']';
            int depth = 0;
            bool in_str = false;
            for (; p < e; ++p) {
                if (in_str) { if (*p == '\\') ++p; else if (*p == '"') in_str = false; continue; }
                if (*p == '"') in_str = true;
                else if (*p == open) ++depth;
                else if (*p == close && --depth == 0) { ++p; return true; }
            }
            return false;
        }
        while (p < e && *p != ',' && *p != '}' && *p != ']') ++p;
        return true;
    }
};

bool parse_header(StFile& f, const char* json, size_t len) {
    Scanner s{json, json + len};
    if (!s.eat('{')) return false;
    if (s.eat('}')) return true;
    do {
        std::string key;
        if (!s.str(key) || !s.eat(':')) return false;
        if (key == "__metadata__") { if (!s.skip_value()) return false; continue; }
        TensorInfo t;
        t.name = key;
        if (!s.eat('{')) return false;
        do {
            std::string k;
            if (!s.str(k) || !s.eat(':')) return false;
            if (k == "dtype") {
                if (!s.str(t.dtype)) return false;
            } else if (k == "shape") {
                if (!s.eat('[')) return false;
                if (!s.eat(']')) {
                    do { int64_t v; if (!s.integer(v)) return false; t.shape.push_back(v); } while (s.eat(','));
                    if (!s.eat(']')) return false;
                }
            } else if (k == "data_offsets") {
                int64_t a, b;
                if (!s.eat('[') || !s.integer(a) || !s.eat(',') || !s.integer(b) || !s.eat(']')) return false;
                t.begin = (uint64_t)a; t.end = (uint64_t)b;
            } else if (!s.skip_value()) {
                return false;
            }
        } while (s.eat(','));
        if (!s.eat('}')) return false;
        f.index[t.name] = (int)f.tensors.size();
        f.tensors.push_back(std::move(t));
    } while (s.eat(','));
    return s.eat('}');
}

struct CopyJob { uint8_t* dst; const uint8_t* src; size_t bytes; };

void run_copies(std::vector<CopyJob>& jobs, int threads) {
    // split big tensors into 8 MB pieces so that the threads stay balanced
    std::vector<CopyJob> pieces;
    constexpr size_t kPiece = 8u << 20;
    for (const CopyJob& j : jobs)
        for (size_t off = 0; off < j.bytes; off += kPiece)
            pieces.push_back({j.dst + off, j.src + off, std::min(kPiece, j.bytes - off)});
    if (threads < 1) threads = 1;
    threads = (int)std::min<size_t>((size_t)threads, std::max<size_t>(pieces.size(), 1));
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= pieces.size()) return;
            std::memcpy(pieces[i].dst, pieces[i].src, pieces[i].bytes);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
}

struct Bucket { int model, batch, seq, graph; };
struct Router {
    std::vector<std::string> models;
    std::vector<Bucket> buckets;
};

}  // namespace

NXD_RT_API const char* nxd_rt_last_error(void) { return g_err; }
NXD_RT_API int nxd_rt_version(void) { return 1; }

// ---------------------------------------------------------------------------------------------------------------------
// safetensors
// ---------------------------------------------------------------------------------------------------------------------
NXD_RT_API void* nxd_st_open(const char* path) {
    StFile* f = new StFile();
    f->fd = ::open(path, O_RDONLY);
    if (f->fd < 0) { fail("cannot open %s", path); delete f; return nullptr; }
    struct stat st;
    if (fstat(f->fd, &st) != 0 || st.st_size < 8) { fail("%s: not a safetensors file", path); ::close(f->fd); delete f; return nullptr; }
    f->size = (size_t)st.st_size;
    f->map = (uint8_t*)mmap(nullptr, f->size, PROT_READ, MAP_PRIVATE, f->fd, 0);
    if (f->map == MAP_FAILED) { fail("%s: mmap failed", path); ::close(f->fd); delete f; return nullptr; }
    uint64_t hlen = 0;
    std::memcpy(&hlen, f->map, 8);
    if (hlen > f->size - 8 || !parse_header(*f, (const char*)f->map + 8, (size_t)hlen)) {
        fail("%s: malformed safetensors header", path);
        munmap(f->map, f->size); ::close(f->fd); delete f;
        return nullptr;
    }
    f->data_off = 8 + (size_t)hlen;
    for (const TensorInfo& t : f->tensors)
        if (t.end < t.begin || f->data_off + t.end > f->size) {
            fail("%s: tensor %s points outside the file", path, t.name.c_str());
            munmap(f->map, f->size); ::close(f->fd); delete f;
            return nullptr;
        }
    madvise(f->map, f->size, MADV_WILLNEED);
    return f;
}

NXD_RT_API void nxd_st_close(void* h) {
    StFile* f = (StFile*)h;
    if (!f) return;
    if (f->map) munmap(f->map, f->size);
    if (f->fd >= 0) ::close(f->fd);
    delete f;
}

NXD_RT_API int nxd_st_num_tensors(void* h) { return h ? (int)((StFile*)h)->tensors.size() : 0; }

// name / dtype are written NUL-terminated (truncated to the capacities); shape gets up to 8 dims.
NXD_RT_API int nxd_st_tensor_info(void* h, int i, char* name, int name_cap, char* dtype, int dtype_cap, int64_t* shape,
                                  int* ndim, uint64_t* nbytes) {
    StFile* f = (StFile*)h;
    if (!f || i < 0 || i >= (int)f->tensors.size()) return fail("tensor index %d out of range", i);
    const TensorInfo& t = f->tensors[i];
    if (name && name_cap > 0) snprintf(name, (size_t)name_cap, "%s", t.name.c_str());
    if (dtype && dtype_cap > 0) snprintf(dtype, (size_t)dtype_cap, "%s", t.dtype.c_str());
    if (ndim) *ndim = (int)t.shape.size();
    if (shape) for (size_t d = 0; d < t.shape.size() && d < 8; ++d) shape[d] = t.shape[d];
    if (nbytes) *nbytes = t.end - t.begin;
    return 0;
}

NXD_RT_API int nxd_st_find(void* h, const char* name) {
    StFile* f = (StFile*)h;
    if (!f) return -1;
    auto it = f->index.find(name);
    return it == f->index.end() ? -1 : it->second;
}

// Copy n tensors (indices idx[i]) into dst[i] (capacity cap[i] bytes each) with `threads` threads.
NXD_RT_API int nxd_st_read_many(void* h, const int* idx, void* const* dst, const uint64_t* cap, int n, int threads) {
    StFile* f = (StFile*)h;
    if (!f) return fail("null handle");
    std::vector<CopyJob> jobs;
    for (int i = 0; i < n; ++i) {
        if (idx[i] < 0 || idx[i] >= (int)f->tensors.size()) return fail("tensor index %d out of range", idx[i]);
        const TensorInfo& t = f->tensors[idx[i]];
        const uint64_t bytes = t.end - t.begin;
        if (cap[i] < bytes) return fail("buffer for %s too small (%llu < %llu)", t.name.c_str(), (unsigned long long)cap[i], (unsigned long long)bytes);
        jobs.push_back({(uint8_t*)dst[i], f->map + f->data_off + t.begin, (size_t)bytes});
    }
    run_copies(jobs, threads);
    return 0;
}

// Rows [r0, r1) of tensor i viewed as [shape[0], rest]: the dim-0 shard of a tensor-parallel rank.
NXD_RT_API int nxd_st_read_rows(void* h, int i, int64_t r0, int64_t r1, void* dst, uint64_t cap, int threads) {
    StFile* f = (StFile*)h;
    if (!f || i < 0 || i >= (int)f->tensors.size()) return fail("tensor index %d out of range", i);
    const TensorInfo& t = f->tensors[i];
    if (t.shape.empty() || r0 < 0 || r1 > t.shape[0] || r0 > r1) return fail("%s: bad row range", t.name.c_str());
    const uint64_t row = t.shape[0] > 0 ? (t.end - t.begin) / (uint64_t)t.shape[0] : 0;
    const uint64_t bytes = row * (uint64_t)(r1 - r0);
    if (cap < bytes) return fail("buffer for %s rows too small", t.name.c_str());
    std::vector<CopyJob> jobs{{(uint8_t*)dst, f->map + f->data_off + t.begin + row * (uint64_t)r0, (size_t)bytes}};
    run_copies(jobs, threads);
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// bucket router
// ---------------------------------------------------------------------------------------------------------------------
NXD_RT_API void* nxd_router_create(void) { return new Router(); }
NXD_RT_API void nxd_router_destroy(void* r) { delete (Router*)r; }

// Register a compiled bucket of model `key` for inputs [batch, seq]; `graph` is an opaque id returned by route().
NXD_RT_API int nxd_router_add(void* r, const char* key, int batch, int seq, int graph) {
    Router* R = (Router*)r;
    if (!R || !key || batch <= 0 || seq <= 0) return fail("router_add: bad arguments");
    int m = -1;
    for (size_t i = 0; i < R->models.size(); ++i) if (R->models[i] == key) m = (int)i;
    if (m < 0) { m = (int)R->models.size(); R->models.emplace_back(key); }
    for (const Bucket& b : R->buckets)
        if (b.model == m && b.batch == batch && b.seq == seq) return fail("router_add: duplicate bucket %s[%d, %d]", key, batch, seq);
    R->buckets.push_back({m, batch, seq, graph});
    return 0;
}

// Exact-shape lookup over all models (the traced-executor contract of the reference): graph id or -1.
NXD_RT_API int nxd_router_route_exact(void* r, int batch, int seq) {
    Router* R = (Router*)r;
    if (!R) return -1;
    for (const Bucket& b : R->buckets) if (b.batch == batch && b.seq == seq) return b.graph;
    return -1;
}

// Smallest bucket of model `key` with bucket.batch >= batch and bucket.seq >= seq (padding router of the runner):
// graph id or -1; the bucket shape is written to out_batch / out_seq.
NXD_RT_API int nxd_router_route_padded(void* r, const char* key, int batch, int seq, int* out_batch, int* out_seq) {
    Router* R = (Router*)r;
    if (!R || !key) return -1;
    const Bucket* best = nullptr;
    for (const Bucket& b : R->buckets) {
        if (R->models[b.model] != key || b.batch < batch || b.seq < seq) continue;
        if (!best || (int64_t)b.batch * b.seq < (int64_t)best->batch * best->seq ||
            ((int64_t)b.batch * b.seq == (int64_t)best->batch * best->seq && b.seq < best->seq))
            best = &b;
    }
    if (!best) return -1;
    if (out_batch) *out_batch = best->batch;
    if (out_seq) *out_seq = best->seq;
    return best->graph;
}
