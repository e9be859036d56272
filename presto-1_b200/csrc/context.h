// context.h — pg_ctx: one device, one stream, stream-ordered device memory, error text, launch counter.
#pragma once
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <exception>
#include <string>
#include <vector>

#include "common.cuh"

namespace pg { struct JitCache; }

struct pg_kernel_record {
    const char* name;
    cudaEvent_t start, stop;
};

struct pg_ctx {
    int device = 0;
    cudaStream_t stream = 0;   // legacy default stream unless pg_ctx_set_stream() is called
    int64_t launches = 0;      // kernels launched through PG_LAUNCH
    int sm_count = 148;        // multiprocessors this context sizes one-wave grids for (sm_total minus the reserved ones)
    int sm_total = 148;
    std::string err;
    // pinned host bounce buffer for small read-backs (flags, counters)
    int64_t* h_scalar = nullptr;  // 64 x int64, pinned
    // optional per-kernel CUDA-event timing (pg_ctx_set_timing): bench.py's live roofline numbers
    bool timing = false;
    std::vector<pg_kernel_record> records;
    std::vector<cudaEvent_t> event_pool;
    pg::JitCache* jit = nullptr;   // run-time specialised kernels (jit.h), created on first use
    // small host outputs are gathered in one device arena and leave with ONE device->pinned copy at the next
    // synchronisation (pg::sync), instead of one blocking copy per column
    struct pending_out { void* host; int64_t off, bytes; };
    void* out_dev = nullptr;
    char* out_pinned = nullptr;
    int64_t out_cap = 0, out_used = 0;
    std::vector<pending_out> out_pending;
    // small host inputs (Trino pages are <= 1 MiB) are staged through one pinned buffer: the copy to the device is then
    // truly asynchronous instead of one blocking pageable copy per array.  Reset at every top-level API call.
    void* in_dev = nullptr;
    char* in_pinned = nullptr;
    int64_t in_cap = 0, in_used = 0;
    int64_t in_flushed = 0;        // staged bytes [in_flushed, in_used) sit in the pinned arena and leave with ONE copy before the next launch
    // landing zone for packed pages that arrive in device memory (pg_hashagg_add_packed_device): pinned, grown on demand
    char* pk_pinned = nullptr;
    int64_t pk_cap = 0;
    // large freed device blocks kept for the next request of about the same size (DevBuf): a join table / aggregation state of
    // several GB is re-created for every page set, and the driver's pool, once other allocations have been carved out of the
    // freed block, stitches a new one together at 5..10 ms per GB.  All blocks belong to this context's stream.
    std::vector<void*> page_batch_pool;     // pinned page accumulators (hashagg.cu: PageBatch) waiting for their next operator
    struct big_block { void* p; int64_t bytes; };
    std::vector<big_block> big_cache;
    int64_t big_cached_bytes = 0;
    size_t big_cache_entries = 32;              // PRESTOGPU_BIG_CACHE_ENTRIES
    int64_t big_cache_bytes = 64LL << 30;       // PRESTOGPU_BIG_CACHE_GB
    int guard_depth = 0;
    int out_live = 0;              // output-arena pieces handed to an OutStage that has not given them back yet
};

namespace pg {
cudaEvent_t timing_event(pg_ctx* ctx);
void jit_cache_free(pg_ctx* ctx);
void page_batch_pool_free(pg_ctx* ctx);     // hashagg.cu
}

namespace pg {

struct Error : public std::exception {
    int32_t code;
    std::string msg;
    Error(int32_t c, std::string m) : code(c), msg(std::move(m)) {}
    const char* what() const noexcept override { return msg.c_str(); }
};

[[noreturn]] static inline void fail(int32_t code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    throw Error(code, buf);
}

#define PG_CUDA(expr)                                                                            \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) ::pg::fail(PG_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

#define PG_CHECK_ARG(cond, ...) do { if (!(cond)) ::pg::fail(PG_E_INVALID_ARGUMENT, __VA_ARGS__); } while (0)
#define PG_CHECK_STATE(cond, ...) do { if (!(cond)) ::pg::fail(PG_E_ILLEGAL_STATE, __VA_ARGS__); } while (0)

// host inputs staged in the pinned arena since the last flush go to the device with one copy (called before anything that may read them)
static inline void flush_staged_inputs(pg_ctx* ctx) {
    if (ctx->in_flushed < ctx->in_used) {
        cudaMemcpyAsync((char*)ctx->in_dev + ctx->in_flushed, ctx->in_pinned + ctx->in_flushed, (size_t)(ctx->in_used - ctx->in_flushed),
                        cudaMemcpyHostToDevice, ctx->stream);
        ctx->in_flushed = ctx->in_used;
    }
}

// kernel<<<grid, block, smem, ctx->stream>>>(...) with launch counting and error check
#define PG_LAUNCH(ctx, kernel, grid, block, smem, ...)                                           \
    do {                                                                                         \
        ::pg::flush_staged_inputs(ctx);                                                          \
        pg_kernel_record _rec;                                                                   \
        if ((ctx)->timing) {                                                                     \
            _rec.name = #kernel; _rec.start = ::pg::timing_event(ctx); _rec.stop = ::pg::timing_event(ctx); \
            cudaEventRecord(_rec.start, (ctx)->stream);                                          \
        }                                                                                        \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                         \
        if ((ctx)->timing) { cudaEventRecord(_rec.stop, (ctx)->stream); (ctx)->records.push_back(_rec); } \
        (ctx)->launches++;                                                                       \
        PG_CUDA(cudaGetLastError());                                                             \
    } while (0)

// Wraps the body of every extern "C" entry point.
template <typename F>
static inline int32_t guard(pg_ctx* ctx, F&& f) {
    struct Depth {
        pg_ctx* c;
        explicit Depth(pg_ctx* ctx) : c(ctx) {
            if (c && c->guard_depth++ == 0) {
                if (c->in_used > 0) {
                    cudaStreamSynchronize(c->stream);   // the previous call's staged copies must have left the pinned buffer
                    c->in_used = 0;
                    c->in_flushed = 0;
                }
                c->out_live = 0;
                if (c->out_pending.empty()) c->out_used = 0;
            }
        }
        ~Depth() { if (c) c->guard_depth--; }
    } depth(ctx);
    try {
        if (ctx) PG_CUDA(cudaSetDevice(ctx->device));
        f();
        return PG_OK;
    } catch (const Error& e) {
        if (ctx) { ctx->err = e.msg; ctx->out_pending.clear(); ctx->out_used = 0; ctx->out_live = 0; }
        return e.code;
    } catch (const std::exception& e) {
        if (ctx) { ctx->err = e.what(); ctx->out_pending.clear(); ctx->out_used = 0; ctx->out_live = 0; }
        return PG_E_INTERNAL;
    }
}

// Stream-ordered device buffer (cudaMallocAsync on the context stream).
static const int64_t BIG_BLOCK = 64LL << 20;        // blocks from this size on go through the context's cache
static inline void big_cache_drop(pg_ctx* c, size_t i) {
    cudaFreeAsync(c->big_cache[i].p, c->stream);
    c->big_cached_bytes -= c->big_cache[i].bytes;
    c->big_cache.erase(c->big_cache.begin() + (long)i);
}
static inline void big_cache_clear(pg_ctx* c) { while (!c->big_cache.empty()) big_cache_drop(c, c->big_cache.size() - 1); }

struct DevBuf {
    pg_ctx* ctx = nullptr;
    void* p = nullptr;
    int64_t bytes = 0;
    int64_t block = 0;       // size of the underlying block when it came from / goes back to the big-block cache (>= bytes)
    DevBuf() {}
    DevBuf(pg_ctx* c, int64_t nbytes) { alloc(c, nbytes); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept { ctx = o.ctx; p = o.p; bytes = o.bytes; block = o.block; o.p = nullptr; o.bytes = 0; o.block = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); ctx = o.ctx; p = o.p; bytes = o.bytes; block = o.block; o.p = nullptr; o.bytes = 0; o.block = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void alloc(pg_ctx* c, int64_t nbytes) {
        release();
        ctx = c; bytes = nbytes; block = 0;
        if (nbytes <= 0) return;
        if (nbytes >= BIG_BLOCK) {
            block = nbytes;
            size_t best = (size_t)-1;        // the smallest cached block that holds the request without wasting more than a quarter
            for (size_t i = 0; i < c->big_cache.size(); i++) {
                const int64_t b = c->big_cache[i].bytes;
                if (b >= nbytes && b <= nbytes + nbytes / 4 && (best == (size_t)-1 || b < c->big_cache[best].bytes)) best = i;
            }
            if (best != (size_t)-1) {
                p = c->big_cache[best].p; block = c->big_cache[best].bytes;
                c->big_cached_bytes -= block;
                c->big_cache.erase(c->big_cache.begin() + (long)best);
                return;
            }
        }
        cudaError_t e = cudaMallocAsync(&p, (size_t)nbytes, c->stream);
        if (e != cudaSuccess && !c->big_cache.empty()) {       // give the cached blocks back and try once more
            cudaGetLastError();
            big_cache_clear(c);
            cudaStreamSynchronize(c->stream);
            e = cudaMallocAsync(&p, (size_t)nbytes, c->stream);
        }
        if (e != cudaSuccess) {
            p = nullptr; bytes = 0; block = 0; cudaGetLastError();
            fail(PG_E_INSUFFICIENT_RESOURCES, "device allocation of %lld bytes failed: %s", (long long)nbytes, cudaGetErrorString(e));
        }
    }
    void release() {
        if (!p) return;
        if (block >= BIG_BLOCK && block <= ctx->big_cache_bytes / 2) {
            ctx->big_cache.push_back(pg_ctx::big_block{p, block});
            ctx->big_cached_bytes += block;
            while (ctx->big_cache.size() > ctx->big_cache_entries || ctx->big_cached_bytes > ctx->big_cache_bytes) big_cache_drop(ctx, 0);   // oldest first
        } else cudaFreeAsync(p, ctx->stream);
        p = nullptr; bytes = 0; block = 0;
    }
    // grow to at least nbytes keeping the first keep_bytes
    void grow(pg_ctx* c, int64_t nbytes, int64_t keep_bytes) {
        if (nbytes <= bytes) return;
        DevBuf n(c, nbytes);
        if (p && keep_bytes > 0) PG_CUDA(cudaMemcpyAsync(n.p, p, (size_t)keep_bytes, cudaMemcpyDeviceToDevice, c->stream));
        *this = std::move(n);
    }
    template <typename T> T* as() const { return (T*)p; }
};

// Fill `bytes` (a multiple of 16, p 16-byte aligned) with the 16-byte pattern {lo, hi} (context.cu: 256-bit stores).  On this GPU
// cudaMemsetAsync and 128-bit stores write 3.9 TB/s, 256-bit stores 7.5 TB/s (tools/fill_ceiling.py), and the tables of this library
// are gigabytes that have to be initialised before every build.
void fill16_async(pg_ctx* ctx, void* p, unsigned long long lo, unsigned long long hi, int64_t bytes);
static inline void memset_async(pg_ctx* ctx, void* p, int v, int64_t bytes) {
    if (bytes <= 0) return;
    if (bytes >= (4 << 20) && (((uintptr_t)p | (uintptr_t)bytes) & 15) == 0) {
        const unsigned long long w = 0x0101010101010101ULL * (unsigned long long)(v & 0xff);
        fill16_async(ctx, p, w, w, bytes);
        return;
    }
    PG_CUDA(cudaMemsetAsync(p, v, (size_t)bytes, ctx->stream));
}
static inline void copy_async(pg_ctx* ctx, void* dst, const void* src, int64_t bytes, cudaMemcpyKind kind) {
    if (kind != cudaMemcpyHostToDevice) flush_staged_inputs(ctx);      // the source may be a staged input
    if (bytes > 0) PG_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, kind, ctx->stream));
}
void flush_host_outputs(pg_ctx* ctx);   // context.cu
static inline void sync(pg_ctx* ctx) {
    if (!ctx->out_pending.empty()) { flush_host_outputs(ctx); return; }
    PG_CUDA(cudaStreamSynchronize(ctx->stream));
}
// carve `bytes` out of the output arena (nullptr when it does not fit: the caller stages the column on its own)
void* out_arena_alloc(pg_ctx* ctx, int64_t bytes, int64_t* off);
// stage `bytes` of host memory through the pinned input arena; returns the device address or nullptr when it does not fit
void* in_arena_stage(pg_ctx* ctx, const void* host, int64_t bytes);

// read back up to 64 int64 scalars from device memory (synchronises the stream)
static inline void read_scalars(pg_ctx* ctx, const void* d, int n, int64_t* out) {
    copy_async(ctx, ctx->h_scalar, d, (int64_t)n * 8, cudaMemcpyDeviceToHost);
    sync(ctx);
    for (int i = 0; i < n; i++) out[i] = ctx->h_scalar[i];
}

// A pg_col imported for the duration of one call: host columns are staged to the device, device columns
// are used in place.
struct ColImport {
    DCol d;
    DevBuf values, nulls, offsets;
    int64_t n = 0;
    int64_t value_bytes = 0;
    const int32_t* host_offsets = nullptr;   // the caller's offsets when the column came from host memory
};
void import_col(pg_ctx* ctx, const pg_col& c, int64_t n, ColImport& out);
// fills value_bytes of device-resident variable width columns (left at -1 by import_col) with one synchronisation
// prefix_rows > 0: prefix_bytes[i] also receives the value bytes of the first prefix_rows rows of column i (0 for fixed width)
void resolve_value_bytes(pg_ctx* ctx, ColImport** cols, int count, int64_t prefix_rows = 0, int64_t* prefix_bytes = nullptr);
// bytes of the value buffer of column c with n rows (reads the last offset for variable width columns)
int64_t col_value_bytes(pg_ctx* ctx, const pg_col& c, int64_t n);

// Output column staging: kernels always write device memory; host outputs are copied back in finish().
struct OutStage {
    pg_ctx* ctx = nullptr;
    pg_col_out* o = nullptr;
    void* d_values = nullptr;
    uint8_t* d_nulls = nullptr;
    int32_t* d_offsets = nullptr;
    DevBuf bv, bn, bo;
    int64_t n = 0;
    int64_t av = -1, an = -1, ao = -1;   // arena offsets (>= 0: the part leaves through the batched copy)
    void* stage(DevBuf& own, int64_t bytes, int64_t& aoff) {
        void* p = out_arena_alloc(ctx, bytes, &aoff);
        if (p) { ctx->out_live++; return p; }
        aoff = -1;
        own.alloc(ctx, bytes > 0 ? bytes : 8);
        return own.p;
    }
    void begin(pg_ctx* c, pg_col_out* out, int64_t rows, int64_t value_bytes) {
        ctx = c; o = out; n = rows;
        PG_CHECK_ARG(out != nullptr, "missing output column");
        PG_CHECK_ARG(out->capacity >= rows, "output column capacity %lld < %lld rows", (long long)out->capacity, (long long)rows);
        bool var = type_is_var(out->type);
        if (var) PG_CHECK_ARG(out->offsets != nullptr && out->values_capacity_bytes >= value_bytes,
                              "variable width output needs offsets and %lld value bytes", (long long)value_bytes);
        if (out->mem == PG_MEM_DEVICE) {
            d_values = out->values; d_nulls = out->nulls; d_offsets = out->offsets;
        } else {
            d_values = stage(bv, value_bytes, av);
            if (out->nulls) d_nulls = (uint8_t*)stage(bn, rows > 0 ? rows : 1, an);
            if (var) d_offsets = (int32_t*)stage(bo, (rows + 1) * 4, ao);
        }
    }
    void finish(int64_t value_bytes) {
        if (o->mem == PG_MEM_DEVICE) return;
        leave(o->values, d_values, value_bytes, av);
        if (o->nulls) leave(o->nulls, d_nulls, n, an);
        if (type_is_var(o->type)) leave(o->offsets, d_offsets, (n + 1) * 4, ao);
    }
    void leave(void* host, const void* dev, int64_t bytes, int64_t aoff) {
        if (aoff >= 0 && ctx->out_live > 0) ctx->out_live--;
        if (bytes <= 0) return;
        if (aoff >= 0) ctx->out_pending.push_back({host, aoff, bytes});
        else copy_async(ctx, host, dev, bytes, cudaMemcpyDeviceToHost);
    }
};

// int32/int64/uint8 result vectors that may live on the host or the device
template <typename T>
struct OutVec {
    pg_ctx* ctx = nullptr;
    T* user = nullptr;
    int32_t mem = PG_MEM_DEVICE;
    T* d = nullptr;
    DevBuf tmp;
    int64_t n = 0;
    void begin(pg_ctx* c, T* out, int32_t out_mem, int64_t count) {
        ctx = c; user = out; mem = out_mem; n = count;
        if (!out) { d = nullptr; return; }
        if (out_mem == PG_MEM_DEVICE) d = out;
        else { tmp.alloc(c, (count > 0 ? count : 1) * (int64_t)sizeof(T)); d = tmp.as<T>(); }
    }
    void finish(int64_t count = -1) {
        if (user && mem != PG_MEM_DEVICE) copy_async(ctx, user, d, (count < 0 ? n : count) * (int64_t)sizeof(T), cudaMemcpyDeviceToHost);
    }
};

static inline int grid_for(pg_ctx* ctx, int64_t n, int block, int per_thread = 1, int max_waves = 8) {
    int64_t g = (n + (int64_t)block * per_thread - 1) / ((int64_t)block * per_thread);
    int64_t cap = (int64_t)ctx->sm_count * max_waves * (2048 / block);
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace pg
