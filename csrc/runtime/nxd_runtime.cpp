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
#include <deque>
#include <map>
#include <mutex>
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

// ---------------------------------------------------------------------------------------------------------------------
// paged KV-cache block manager
//
// The device cache of a layer is a pool of pages [num_pages, kv_heads, page_size, head_dim]; a sequence (cache line) owns an
// ordered list of pages, the block table row the attention kernels index with.  The pool may be smaller than
// lines * max_pages: pages are handed out as sequences grow and return to the free list when a line is released.  Pages
// are reference counted so that a line can be forked (speculation branches, shared prompt prefixes): full pages are
// shared, the partially filled last page is copied by the caller (the manager reports which page to copy where).
// ---------------------------------------------------------------------------------------------------------------------
struct KvPool {
    int num_pages, page_size, max_lines, max_pages;
    std::vector<int> free_list;                  // LIFO: recently freed pages are reused first (L2-warm)
    std::vector<int> refcnt;                     // per page
    std::vector<std::vector<int>> pages;         // per line
    std::vector<int64_t> tokens;                 // per line: tokens reserved so far
    std::mutex mu;
};

NXD_RT_API void* nxd_kv_create(int num_pages, int page_size, int max_lines, int max_pages_per_line) {
    if (num_pages <= 0 || page_size <= 0 || max_lines <= 0 || max_pages_per_line <= 0) {
        fail("kv_create: bad arguments");
        return nullptr;
    }
    KvPool* p = new KvPool();
    p->num_pages = num_pages; p->page_size = page_size; p->max_lines = max_lines; p->max_pages = max_pages_per_line;
    p->refcnt.assign(num_pages, 0);
    p->pages.resize(max_lines);
    p->tokens.assign(max_lines, 0);
    p->free_list.reserve(num_pages);
    for (int i = num_pages - 1; i >= 0; --i) p->free_list.push_back(i);     // page 0 is handed out first
    return p;
}

NXD_RT_API void nxd_kv_destroy(void* h) { delete (KvPool*)h; }

NXD_RT_API int nxd_kv_num_free(void* h) { return h ? (int)((KvPool*)h)->free_list.size() : -1; }

NXD_RT_API int64_t nxd_kv_line_tokens(void* h, int line) {
    KvPool* p = (KvPool*)h;
    return (p && line >= 0 && line < p->max_lines) ? p->tokens[line] : -1;
}

// Pages line `line` needs IN ADDITION to what it holds to store `n_tokens` tokens in total.
NXD_RT_API int nxd_kv_pages_needed(void* h, int line, int64_t n_tokens) {
    KvPool* p = (KvPool*)h;
    if (!p || line < 0 || line >= p->max_lines) return -1;
    int64_t want = (n_tokens + p->page_size - 1) / p->page_size;
    int64_t have = (int64_t)p->pages[line].size();
    return (int)(want > have ? want - have : 0);
}

// Make room for `n_tokens` tokens in total on `line`.  Returns the number of pages newly allocated, -1 when the pool (or
// the per-line limit) cannot hold them -- in that case nothing is allocated.
NXD_RT_API int nxd_kv_reserve(void* h, int line, int64_t n_tokens) {
    KvPool* p = (KvPool*)h;
    if (!p || line < 0 || line >= p->max_lines || n_tokens < 0) return fail("kv_reserve: bad arguments");
    std::lock_guard<std::mutex> g(p->mu);
    int64_t want = (n_tokens + p->page_size - 1) / p->page_size;
    if (want > p->max_pages) return fail("kv_reserve: %lld tokens exceed the per-line limit of %d pages", (long long)n_tokens, p->max_pages);
    int64_t have = (int64_t)p->pages[line].size();
    int need = (int)(want > have ? want - have : 0);
    if (need > (int)p->free_list.size()) return fail("kv_reserve: out of pages (%d needed, %d free)", need, (int)p->free_list.size());
    for (int i = 0; i < need; ++i) {
        int pg = p->free_list.back();
        p->free_list.pop_back();
        p->refcnt[pg] = 1;
        p->pages[line].push_back(pg);
    }
    if (n_tokens > p->tokens[line]) p->tokens[line] = n_tokens;
    return need;
}

// Release every page of `line` (pages shared with a fork stay alive until their last owner lets go).  Returns pages freed.
NXD_RT_API int nxd_kv_free_line(void* h, int line) {
    KvPool* p = (KvPool*)h;
    if (!p || line < 0 || line >= p->max_lines) return fail("kv_free_line: bad line");
    std::lock_guard<std::mutex> g(p->mu);
    int freed = 0;
    for (auto it = p->pages[line].rbegin(); it != p->pages[line].rend(); ++it)
        if (--p->refcnt[*it] == 0) { p->free_list.push_back(*it); ++freed; }
    p->pages[line].clear();
    p->tokens[line] = 0;
    return freed;
}

// Drop the pages of `line` beyond `n_tokens` (speculation rollback).  Returns pages freed.
NXD_RT_API int nxd_kv_truncate(void* h, int line, int64_t n_tokens) {
    KvPool* p = (KvPool*)h;
    if (!p || line < 0 || line >= p->max_lines || n_tokens < 0) return fail("kv_truncate: bad arguments");
    std::lock_guard<std::mutex> g(p->mu);
    size_t keep = (size_t)((n_tokens + p->page_size - 1) / p->page_size);
    int freed = 0;
    while (p->pages[line].size() > keep) {
        int pg = p->pages[line].back();
        p->pages[line].pop_back();
        if (--p->refcnt[pg] == 0) { p->free_list.push_back(pg); ++freed; }
    }
    if (n_tokens < p->tokens[line]) p->tokens[line] = n_tokens;
    return freed;
}

// dst becomes a copy-on-write fork of src: full pages are shared (reference counted); when the last page of src is only
// partially filled, dst gets a fresh page for it and (*copy_src, *copy_dst) tell the caller which page to copy on the device
// (both -1 when no copy is needed).  Returns 0, or -1 when the pool is exhausted.
NXD_RT_API int nxd_kv_fork(void* h, int src, int dst, int* copy_src, int* copy_dst) {
    KvPool* p = (KvPool*)h;
    if (!p || src < 0 || dst < 0 || src >= p->max_lines || dst >= p->max_lines || src == dst) return fail("kv_fork: bad lines");
    std::lock_guard<std::mutex> g(p->mu);
    if (!p->pages[dst].empty()) return fail("kv_fork: destination line %d is in use", dst);
    if (copy_src) *copy_src = -1;
    if (copy_dst) *copy_dst = -1;
    const std::vector<int>& sp = p->pages[src];
    bool partial = !sp.empty() && (p->tokens[src] % p->page_size) != 0;
    if (partial && p->free_list.empty()) return fail("kv_fork: out of pages");
    size_t shared = partial ? sp.size() - 1 : sp.size();
    for (size_t i = 0; i < shared; ++i) { ++p->refcnt[sp[i]]; p->pages[dst].push_back(sp[i]); }
    if (partial) {
        int pg = p->free_list.back();
        p->free_list.pop_back();
        p->refcnt[pg] = 1;
        p->pages[dst].push_back(pg);
        if (copy_src) *copy_src = sp.back();
        if (copy_dst) *copy_dst = pg;
    }
    p->tokens[dst] = p->tokens[src];
    return 0;
}

// Is the page that position `pos` of `line` falls on shared with another line?  (A writer must un-share it first.)
NXD_RT_API int nxd_kv_page_shared(void* h, int line, int64_t pos) {
    KvPool* p = (KvPool*)h;
    if (!p || line < 0 || line >= p->max_lines || pos < 0) return -1;
    size_t i = (size_t)(pos / p->page_size);
    if (i >= p->pages[line].size()) return 0;
    return p->refcnt[p->pages[line][i]] > 1 ? 1 : 0;
}

// Replace the shared page under position `pos` of `line` by a private one; (*copy_src, *copy_dst) as in nxd_kv_fork.
NXD_RT_API int nxd_kv_unshare(void* h, int line, int64_t pos, int* copy_src, int* copy_dst) {
    KvPool* p = (KvPool*)h;
    if (!p || line < 0 || line >= p->max_lines || pos < 0) return fail("kv_unshare: bad arguments");
    std::lock_guard<std::mutex> g(p->mu);
    if (copy_src) *copy_src = -1;
    if (copy_dst) *copy_dst = -1;
    size_t i = (size_t)(pos / p->page_size);
    if (i >= p->pages[line].size()) return 0;
    int old = p->pages[line][i];
    if (p->refcnt[old] <= 1) return 0;
    if (p->free_list.empty()) return fail("kv_unshare: out of pages");
    int pg = p->free_list.back();
    p->free_list.pop_back();
    p->refcnt[pg] = 1;
    --p->refcnt[old];
    p->pages[line][i] = pg;
    if (copy_src) *copy_src = old;
    if (copy_dst) *copy_dst = pg;
    return 1;
}

// Block-table row of `line` into out[max_pages]; unused slots get `fill` (a valid page id keeps kernels in bounds).
NXD_RT_API int nxd_kv_block_row(void* h, int line, int32_t* out, int fill) {
    KvPool* p = (KvPool*)h;
    if (!p || !out || line < 0 || line >= p->max_lines) return fail("kv_block_row: bad arguments");
    const std::vector<int>& pg = p->pages[line];
    for (int i = 0; i < p->max_pages; ++i) out[i] = i < (int)pg.size() ? pg[i] : fill;
    return (int)pg.size();
}

// ---------------------------------------------------------------------------------------------------------------------
// continuous-batching scheduler (host side): which waiting request is admitted to which cache line, given the pages free
// ---------------------------------------------------------------------------------------------------------------------
struct SchedReq { int64_t id; int prompt_len; int max_new; int line; int generated; };
struct Scheduler {
    KvPool* pool;
    int max_running;
    std::deque<SchedReq> waiting;
    std::vector<SchedReq> running;
    std::vector<int> free_lines;
};

NXD_RT_API void* nxd_sched_create(void* kv_pool, int max_running) {
    KvPool* p = (KvPool*)kv_pool;
    if (!p || max_running <= 0 || max_running > p->max_lines) { fail("sched_create: bad arguments"); return nullptr; }
    Scheduler* s = new Scheduler();
    s->pool = p; s->max_running = max_running;
    for (int i = max_running - 1; i >= 0; --i) s->free_lines.push_back(i);
    return s;
}
NXD_RT_API void nxd_sched_destroy(void* h) { delete (Scheduler*)h; }

NXD_RT_API int nxd_sched_submit(void* h, int64_t id, int prompt_len, int max_new_tokens) {
    Scheduler* s = (Scheduler*)h;
    if (!s || prompt_len <= 0 || max_new_tokens <= 0) return fail("sched_submit: bad arguments");
    if ((int64_t)prompt_len + max_new_tokens > (int64_t)s->pool->max_pages * s->pool->page_size)
        return fail("sched_submit: request %lld does not fit a cache line", (long long)id);
    s->waiting.push_back({id, prompt_len, max_new_tokens, -1, 0});
    return 0;
}

// Admit waiting requests in arrival order while a line is free and the pool holds their prompt plus one page of head room
// per running sequence (so that admitted work can always take its next decode step).  Writes (id, line) pairs; returns count.
NXD_RT_API int nxd_sched_admit(void* h, int64_t* ids, int* lines, int cap) {
    Scheduler* s = (Scheduler*)h;
    if (!s) return -1;
    int n = 0;
    while (n < cap && !s->waiting.empty() && !s->free_lines.empty()) {
        SchedReq r = s->waiting.front();
        int prompt_pages = (r.prompt_len + 1 + s->pool->page_size - 1) / s->pool->page_size;     // + the first new token
        int headroom = (int)s->running.size();
        if (prompt_pages + headroom > (int)s->pool->free_list.size()) break;
        int line = s->free_lines.back();
        if (nxd_kv_reserve(s->pool, line, r.prompt_len + 1) < 0) break;
        s->free_lines.pop_back();
        s->waiting.pop_front();
        r.line = line;
        s->running.push_back(r);
        if (ids) ids[n] = r.id;
        if (lines) lines[n] = line;
        ++n;
    }
    return n;
}

// One decode step was taken by every running request: grow their reservations by one token.  Requests that reached
// max_new_tokens (or whose id is in done_ids) are retired and their pages freed.  A request that cannot get its next page is
// preempted: pages freed, request put back at the FRONT of the waiting queue (it will re-run its prompt).  Writes the ids of
// retired requests to out_done and of preempted ones to out_preempted; returns the number still running.
NXD_RT_API int nxd_sched_step(void* h, const int64_t* done_ids, int n_done, int64_t* out_done, int* n_out_done,
                              int64_t* out_preempted, int* n_out_preempted, int cap) {
    Scheduler* s = (Scheduler*)h;
    if (!s) return -1;
    int nd = 0, np = 0;
    std::vector<SchedReq> keep;
    for (SchedReq& r : s->running) {
        ++r.generated;
        bool done = r.generated >= r.max_new;
        for (int i = 0; i < n_done && !done; ++i) done = done_ids[i] == r.id;
        if (done) {
            nxd_kv_free_line(s->pool, r.line);
            s->free_lines.push_back(r.line);
            if (out_done && nd < cap) out_done[nd] = r.id;
            ++nd;
            continue;
        }
        keep.push_back(r);
    }
    s->running.clear();
    // grow oldest first; the youngest sequences are the ones preempted when the pool runs dry
    for (SchedReq& r : keep) {
        if (nxd_kv_reserve(s->pool, r.line, (int64_t)r.prompt_len + r.generated + 1) < 0) {
            nxd_kv_free_line(s->pool, r.line);
            s->free_lines.push_back(r.line);
            if (out_preempted && np < cap) out_preempted[np] = r.id;
            ++np;
            SchedReq w = r;
            w.line = -1; w.generated = 0;
            s->waiting.push_front(w);
            continue;
        }
        s->running.push_back(r);
    }
    if (n_out_done) *n_out_done = nd;
    if (n_out_preempted) *n_out_preempted = np;
    return (int)s->running.size();
}

NXD_RT_API int nxd_sched_num_waiting(void* h) { return h ? (int)((Scheduler*)h)->waiting.size() : -1; }
NXD_RT_API int nxd_sched_num_running(void* h) { return h ? (int)((Scheduler*)h)->running.size() : -1; }
