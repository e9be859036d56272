// context.cu — context lifetime, column import (host Blocks -> device-resident column buffers).
#include <stdlib.h>

#include <algorithm>

#include "context.h"

namespace pg {

// 256-bit stores (STG.E.ENL2.256): one instruction moves 32 bytes per lane, and the store path of an SM is bound by instructions, not bytes
__device__ __forceinline__ void st256(void* p, unsigned long long a, unsigned long long b) {
    asm volatile("st.global.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(p), "l"(a), "l"(b), "l"(a), "l"(b) : "memory");
}
__global__ void __launch_bounds__(256) fill16_kernel(unsigned long long* body, int64_t n32, unsigned long long* head, unsigned long long* tail,
                                                     unsigned long long lo, unsigned long long hi) {
    const int64_t base = (int64_t)blockIdx.x * 1024 + threadIdx.x;     // a block fills 32 KB in a row
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int64_t u = base + k * 256;
        if (u < n32) st256(body + 4 * u, lo, hi);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (head) { head[0] = lo; head[1] = hi; }
        if (tail) { tail[0] = lo; tail[1] = hi; }
    }
}
void fill16_async(pg_ctx* ctx, void* p, unsigned long long lo, unsigned long long hi, int64_t bytes) {
    if (bytes <= 0) return;
    PG_CHECK_STATE((((uintptr_t)p | (uintptr_t)bytes) & 15) == 0, "fill16: pointer and size must be multiples of 16");
    unsigned long long* q = (unsigned long long*)p;
    unsigned long long* head = nullptr;
    if ((uintptr_t)q & 31) { head = q; q += 2; bytes -= 16; }
    const int64_t n32 = bytes / 32;
    unsigned long long* tail = (bytes & 31) ? q + 4 * n32 : nullptr;
    const int64_t blocks = std::max<int64_t>(1, (n32 + 1023) / 1024);
    PG_LAUNCH(ctx, fill16_kernel, (int)blocks, 256, 0, q, n32, head, tail, lo, hi);
}

int64_t col_value_bytes(pg_ctx* ctx, const pg_col& c, int64_t n) {
    if (!type_is_var(c.type)) return n * type_width(c.type);
    if (n == 0) return 0;
    PG_CHECK_ARG(c.offsets != nullptr, "variable width column without offsets");
    int32_t last = 0, first = 0;
    if (c.mem == PG_MEM_DEVICE) {
        copy_async(ctx, ctx->h_scalar, c.offsets + n, 4, cudaMemcpyDeviceToHost);
        copy_async(ctx, (char*)ctx->h_scalar + 8, c.offsets, 4, cudaMemcpyDeviceToHost);
        sync(ctx);
        memcpy(&last, ctx->h_scalar, 4);
        memcpy(&first, (char*)ctx->h_scalar + 8, 4);
    } else {
        last = c.offsets[n]; first = c.offsets[0];
    }
    (void)first;
    return last;  // offsets index into `values` from its start
}

void import_col(pg_ctx* ctx, const pg_col& c, int64_t n, ColImport& out) {
    PG_CHECK_ARG(type_valid(c.type), "unknown column type %d", c.type);
    PG_CHECK_ARG(n >= 0, "negative row count");
    PG_CHECK_ARG(n == 0 || c.values != nullptr || type_is_var(c.type), "column without values");
    out.n = n;
    out.d.type = c.type; out.d.pad = 0;
    if (c.mem == PG_MEM_DEVICE) {
        // in place; the byte size of a variable width value buffer sits in device memory (offsets[n]) and is only
        // fetched by the callers that need it (resolve_value_bytes), one synchronisation for all columns
        out.value_bytes = type_is_var(c.type) ? -1 : n * type_width(c.type);
        out.d.values = c.values; out.d.nulls = c.nulls; out.d.offsets = c.offsets;
        return;
    }
    out.value_bytes = col_value_bytes(ctx, c, n);
    out.host_offsets = type_is_var(c.type) ? c.offsets : nullptr;
    PG_CHECK_ARG(c.mem == PG_MEM_HOST || c.mem == PG_MEM_HOST_PINNED, "bad mem kind %d", c.mem);
    // small arrays go through the pinned arena (asynchronous), large ones straight from the caller's memory
    auto stage = [&](DevBuf& own, const void* host, int64_t bytes) -> void* {
        if (void* p = in_arena_stage(ctx, host, bytes)) return p;
        own.alloc(ctx, bytes > 0 ? bytes : 8);
        copy_async(ctx, own.p, host, bytes, cudaMemcpyHostToDevice);
        return own.p;
    };
    out.d.values = stage(out.values, c.values, out.value_bytes);
    out.d.nulls = nullptr; out.d.offsets = nullptr;
    if (c.nulls) out.d.nulls = (const uint8_t*)stage(out.nulls, c.nulls, n);
    if (type_is_var(c.type)) {
        if (c.offsets) out.d.offsets = (const int32_t*)stage(out.offsets, c.offsets, (n + 1) * 4);
        else {
            out.offsets.alloc(ctx, (n + 1) * 4);
            memset_async(ctx, out.offsets.p, 0, (n + 1) * 4);
            out.d.offsets = out.offsets.as<int32_t>();
        }
    }
}

void resolve_value_bytes(pg_ctx* ctx, ColImport** cols, int count, int64_t prefix_rows, int64_t* prefix_bytes) {
    int pending[32], np = 0;
    const bool want_prefix = prefix_rows > 0 && prefix_bytes != nullptr;
    for (int i = 0; i < count && np < 32; i++) {
        ColImport* c = cols[i];
        if (want_prefix) prefix_bytes[i] = 0;
        if (!type_is_var(c->d.type)) continue;
        if (c->n == 0) { c->value_bytes = 0; continue; }
        const int64_t pr = prefix_rows < c->n ? prefix_rows : c->n;
        if (c->value_bytes >= 0 && (!want_prefix || c->host_offsets)) {
            if (want_prefix && c->host_offsets) prefix_bytes[i] = c->host_offsets[pr];
            else if (want_prefix) prefix_bytes[i] = c->value_bytes;
            continue;
        }
        copy_async(ctx, (char*)ctx->h_scalar + 8 * np, c->d.offsets + c->n, 4, cudaMemcpyDeviceToHost);
        if (want_prefix) copy_async(ctx, (char*)ctx->h_scalar + 8 * np + 4, c->d.offsets + pr, 4, cudaMemcpyDeviceToHost);
        pending[np++] = i;
    }
    if (np == 0) return;
    sync(ctx);
    for (int j = 0; j < np; j++) {
        int32_t v[2];
        memcpy(v, (char*)ctx->h_scalar + 8 * j, 8);
        cols[pending[j]]->value_bytes = v[0];
        if (want_prefix) prefix_bytes[pending[j]] = v[1];
    }
}

static const int64_t OUT_ARENA_BYTES = 4 << 20, OUT_ARENA_MAX_PIECE = 512 << 10;

void* out_arena_alloc(pg_ctx* ctx, int64_t bytes, int64_t* off) {
    if (bytes > OUT_ARENA_MAX_PIECE) return nullptr;
    if (!ctx->out_dev) {
        if (cudaMalloc(&ctx->out_dev, OUT_ARENA_BYTES) != cudaSuccess) { cudaGetLastError(); ctx->out_dev = nullptr; return nullptr; }
        if (cudaMallocHost((void**)&ctx->out_pinned, OUT_ARENA_BYTES) != cudaSuccess) {
            cudaGetLastError(); cudaFree(ctx->out_dev); ctx->out_dev = nullptr; return nullptr;
        }
        ctx->out_cap = OUT_ARENA_BYTES;
    }
    int64_t need = (bytes + 255) & ~255LL;
    if (need == 0) need = 256;
    if (ctx->out_used + need > ctx->out_cap) return nullptr;
    *off = ctx->out_used;
    ctx->out_used += need;
    return (char*)ctx->out_dev + *off;
}

void* in_arena_stage(pg_ctx* ctx, const void* host, int64_t bytes) {
    const int64_t ARENA = 4 << 20, MAX_PIECE = 256 << 10;
    if (bytes <= 0 || bytes > MAX_PIECE) return nullptr;
    if (!ctx->in_dev) {
        if (cudaMalloc(&ctx->in_dev, ARENA) != cudaSuccess) { cudaGetLastError(); ctx->in_dev = nullptr; return nullptr; }
        if (cudaMallocHost((void**)&ctx->in_pinned, ARENA) != cudaSuccess) {
            cudaGetLastError(); cudaFree(ctx->in_dev); ctx->in_dev = nullptr; return nullptr;
        }
        ctx->in_cap = ARENA;
    }
    const int64_t need = (bytes + 255) & ~255LL;
    if (ctx->in_used + need > ctx->in_cap) return nullptr;
    const int64_t off = ctx->in_used;
    ctx->in_used += need;
    memcpy(ctx->in_pinned + off, host, (size_t)bytes);      // leaves for the device with the next flush_staged_inputs (one copy for all pieces)
    return (char*)ctx->in_dev + off;
}

void flush_host_outputs(pg_ctx* ctx) {
    cudaError_t e = cudaMemcpyAsync(ctx->out_pinned, ctx->out_dev, (size_t)ctx->out_used, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    std::vector<pg_ctx::pending_out> pend;
    pend.swap(ctx->out_pending);
    if (ctx->out_live == 0) ctx->out_used = 0;   // pieces still being filled keep their place
    if (e != cudaSuccess) fail(PG_E_CUDA, "flushing host outputs failed: %s", cudaGetErrorString(e));
    for (const pg_ctx::pending_out& p : pend) memcpy(p.host, ctx->out_pinned + p.off, (size_t)p.bytes);
}

cudaEvent_t timing_event(pg_ctx* ctx) {
    if (!ctx->event_pool.empty()) { cudaEvent_t e = ctx->event_pool.back(); ctx->event_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

}  // namespace pg

using namespace pg;

extern "C" {

int32_t pg_ctx_create(int32_t device, pg_ctx** out) {
    if (!out) return PG_E_INVALID_ARGUMENT;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) { cudaGetLastError(); return PG_E_CUDA; }   // no CPU fallback: fail loudly
    if (device < 0 || device >= count) return PG_E_INVALID_ARGUMENT;
    pg_ctx* ctx = new pg_ctx();
    ctx->device = device;
    int32_t rc = guard(ctx, [&] {
        cudaDeviceProp prop;
        PG_CUDA(cudaGetDeviceProperties(&prop, device));
        ctx->sm_count = ctx->sm_total = prop.multiProcessorCount;
        PG_CUDA(cudaMallocHost((void**)&ctx->h_scalar, 64 * sizeof(int64_t)));
        // keep freed blocks cached in the default pool: per-page scratch is reused, not returned to the OS
        cudaMemPool_t pool;
        PG_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        uint64_t thresh = UINT64_MAX;
        PG_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
        // L2 fetch granularity on a miss (32 / 64 / 128 bytes; the driver's default is 64).  Hash probes and gathers want one 32-byte
        // sector per random access; PRESTOGPU_L2_FETCH overrides (0 = leave the device setting alone).
        // blocks of 64 MB and more are kept by the context when freed (context.h: DevBuf) instead of going back to the pool
        if (const char* e = getenv("PRESTOGPU_BIG_CACHE_ENTRIES")) ctx->big_cache_entries = (size_t)std::max(0, atoi(e));
        if (const char* e = getenv("PRESTOGPU_BIG_CACHE_GB")) ctx->big_cache_bytes = (int64_t)std::max(0, atoi(e)) << 30;
        int fetch = 0;
        if (const char* e = getenv("PRESTOGPU_L2_FETCH")) fetch = atoi(e);
        if (fetch == 32 || fetch == 64 || fetch == 128) {
            if (cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)fetch) != cudaSuccess) cudaGetLastError();
        }
    });
    if (rc != PG_OK) { delete ctx; return rc; }
    *out = ctx;
    return PG_OK;
}

int32_t pg_ctx_destroy(pg_ctx* ctx) {
    if (!ctx) return PG_OK;
    cudaSetDevice(ctx->device);
    pg::page_batch_pool_free(ctx);
    pg::big_cache_clear(ctx);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->h_scalar) cudaFreeHost(ctx->h_scalar);
    if (ctx->out_dev) cudaFree(ctx->out_dev);
    if (ctx->out_pinned) cudaFreeHost(ctx->out_pinned);
    if (ctx->in_dev) cudaFree(ctx->in_dev);
    if (ctx->in_pinned) cudaFreeHost(ctx->in_pinned);
    if (ctx->pk_pinned) cudaFreeHost(ctx->pk_pinned);
    for (const pg_kernel_record& r : ctx->records) { cudaEventDestroy(r.start); cudaEventDestroy(r.stop); }
    for (cudaEvent_t e : ctx->event_pool) cudaEventDestroy(e);
    pg::jit_cache_free(ctx);
    delete ctx;
    return PG_OK;
}

const char* pg_last_error(pg_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int32_t pg_ctx_set_stream(pg_ctx* ctx, void* cuda_stream) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    if (ctx->stream != (cudaStream_t)cuda_stream) { cudaSetDevice(ctx->device); pg::big_cache_clear(ctx); }   // cached blocks are ordered on the old stream
    ctx->stream = (cudaStream_t)cuda_stream;
    return PG_OK;
}

int32_t pg_ctx_reserve_sms(pg_ctx* ctx, int32_t sms) {
    if (!ctx || sms < 0 || sms >= ctx->sm_total) return PG_E_INVALID_ARGUMENT;
    ctx->sm_count = ctx->sm_total - sms;
    return PG_OK;
}

int32_t pg_ctx_synchronize(pg_ctx* ctx) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] { sync(ctx); });
}

int64_t pg_ctx_launch_count(pg_ctx* ctx) { return ctx ? ctx->launches : 0; }

int32_t pg_ctx_set_timing(pg_ctx* ctx, int32_t enabled) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    ctx->timing = enabled != 0;
    return PG_OK;
}

/* Sums the recorded launches whose kernel name contains `name_substr` (NULL = all), then clears them when `reset`. */
int32_t pg_ctx_kernel_time(pg_ctx* ctx, const char* name_substr, int32_t reset, double* total_ms, int64_t* launches) {
    if (!ctx) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        sync(ctx);
        double ms = 0; int64_t cnt = 0;
        for (const pg_kernel_record& r : ctx->records) {
            if (name_substr && !strstr(r.name, name_substr)) continue;
            float t = 0;
            if (cudaEventElapsedTime(&t, r.start, r.stop) == cudaSuccess) { ms += t; cnt++; }
        }
        if (total_ms) *total_ms = ms;
        if (launches) *launches = cnt;
        if (reset) {
            for (const pg_kernel_record& r : ctx->records) { ctx->event_pool.push_back(r.start); ctx->event_pool.push_back(r.stop); }
            ctx->records.clear();
        }
    });
}

/* Writes "name=ms/launches;..." for every kernel recorded since the last reset (for bench.py's launch breakdown). */
int32_t pg_ctx_kernel_report(pg_ctx* ctx, char* buf, int64_t cap) {
    if (!ctx || !buf || cap <= 0) return PG_E_INVALID_ARGUMENT;
    return guard(ctx, [&] {
        sync(ctx);
        std::vector<std::string> names; std::vector<double> ms; std::vector<int64_t> cnt;
        for (const pg_kernel_record& r : ctx->records) {
            float t = 0;
            if (cudaEventElapsedTime(&t, r.start, r.stop) != cudaSuccess) continue;
            size_t k = 0;
            for (; k < names.size(); k++) if (names[k] == r.name) break;
            if (k == names.size()) { names.push_back(r.name); ms.push_back(0); cnt.push_back(0); }
            ms[k] += t; cnt[k]++;
        }
        std::string out;
        char line[256];
        for (size_t k = 0; k < names.size(); k++) {
            snprintf(line, sizeof(line), "%s=%.6f/%lld;", names[k].c_str(), ms[k], (long long)cnt[k]);
            out += line;
        }
        snprintf(buf, (size_t)cap, "%s", out.c_str());
    });
}

const char* pg_version(void) { return "prestogpu 0.1 (sm_100a)"; }

}  // extern "C"
