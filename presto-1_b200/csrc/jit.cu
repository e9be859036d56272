// jit.cu — NVRTC loader, per-context kernel cache, launch helper (see jit.h).
#include "jit.h"

#include <dlfcn.h>
#include <stdlib.h>

#include <chrono>
#include <mutex>

namespace pg {

// ---- the handful of NVRTC entry points we need, resolved with dlsym (nvrtc.h is not required at build time)
typedef struct _nvrtcProgram* nvrtcProgram_t;
typedef int (*fn_nvrtcCreateProgram)(nvrtcProgram_t*, const char*, const char*, int, const char* const*, const char* const*);
typedef int (*fn_nvrtcCompileProgram)(nvrtcProgram_t, int, const char* const*);
typedef int (*fn_nvrtcDestroyProgram)(nvrtcProgram_t*);
typedef int (*fn_nvrtcGetSize)(nvrtcProgram_t, size_t*);
typedef int (*fn_nvrtcGetData)(nvrtcProgram_t, char*);
typedef const char* (*fn_nvrtcGetErrorString)(int);

struct Nvrtc {
    void* handle = nullptr;
    fn_nvrtcCreateProgram create = nullptr;
    fn_nvrtcCompileProgram compile = nullptr;
    fn_nvrtcDestroyProgram destroy = nullptr;
    fn_nvrtcGetSize cubin_size = nullptr, log_size = nullptr;
    fn_nvrtcGetData cubin = nullptr, log = nullptr;
    fn_nvrtcGetErrorString error_string = nullptr;
    std::string why_not;
};

static Nvrtc* nvrtc() {
    static Nvrtc lib;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* env = getenv("PRESTOGPU_JIT");
        if (env && env[0] == '0') { lib.why_not = "disabled by PRESTOGPU_JIT=0"; return; }
        const char* names[] = {getenv("PRESTOGPU_NVRTC"), "libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so",
                               "/usr/local/cuda/lib64/libnvrtc.so"};
        for (const char* name : names) {
            if (!name || !name[0]) continue;
            lib.handle = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (lib.handle) break;
        }
        if (!lib.handle) { lib.why_not = "libnvrtc.so.12 not found (set PRESTOGPU_NVRTC to its path)"; return; }
        lib.create = (fn_nvrtcCreateProgram)dlsym(lib.handle, "nvrtcCreateProgram");
        lib.compile = (fn_nvrtcCompileProgram)dlsym(lib.handle, "nvrtcCompileProgram");
        lib.destroy = (fn_nvrtcDestroyProgram)dlsym(lib.handle, "nvrtcDestroyProgram");
        lib.cubin_size = (fn_nvrtcGetSize)dlsym(lib.handle, "nvrtcGetCUBINSize");
        lib.cubin = (fn_nvrtcGetData)dlsym(lib.handle, "nvrtcGetCUBIN");
        lib.log_size = (fn_nvrtcGetSize)dlsym(lib.handle, "nvrtcGetProgramLogSize");
        lib.log = (fn_nvrtcGetData)dlsym(lib.handle, "nvrtcGetProgramLog");
        lib.error_string = (fn_nvrtcGetErrorString)dlsym(lib.handle, "nvrtcGetErrorString");
        if (!lib.create || !lib.compile || !lib.destroy || !lib.cubin_size || !lib.cubin || !lib.log_size || !lib.log) {
            lib.why_not = "libnvrtc lacks the cubin API (needs CUDA 11.1+)";
            lib.handle = nullptr;
        }
    });
    return &lib;
}

bool jit_available(std::string* why_not) {
    Nvrtc* n = nvrtc();
    if (!n->handle && why_not) *why_not = n->why_not;
    return n->handle != nullptr;
}

JitCache::~JitCache() {
    for (auto& kv : kernels)
        if (kv.second.lib) cudaLibraryUnload(kv.second.lib);
}

// compile `source` to an sm_100a cubin; returns false with the compiler log in `err`
bool jit_compile_cubin(const std::string& source, std::string& cubin, std::string& err) {
    Nvrtc* n = nvrtc();
    if (!n->handle) { err = n->why_not; return false; }
    nvrtcProgram_t prog = nullptr;
    int rc = n->create(&prog, source.c_str(), "pg_jit.cu", 0, nullptr, nullptr);
    if (rc != 0) { err = std::string("nvrtcCreateProgram: ") + (n->error_string ? n->error_string(rc) : "error"); return false; }
    const char* opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "-lineinfo", "-default-device"};
    rc = n->compile(prog, 4, opts);
    if (rc != 0) {
        size_t ls = 0;
        n->log_size(prog, &ls);
        std::string log(ls, '\0');
        if (ls) n->log(prog, &log[0]);
        err = std::string("nvrtcCompileProgram: ") + (n->error_string ? n->error_string(rc) : "error") + "\n" + log;
        n->destroy(&prog);
        return false;
    }
    size_t cs = 0;
    n->cubin_size(prog, &cs);
    cubin.assign(cs, '\0');
    if (cs) n->cubin(prog, &cubin[0]);
    n->destroy(&prog);
    if (cs == 0) { err = "NVRTC produced no cubin"; return false; }
    return true;
}

JitKernel* jit_get(pg_ctx* ctx, const std::string& signature, const std::string& source, const char* entry) {
    if (!ctx->jit) ctx->jit = new JitCache();
    JitCache& cache = *ctx->jit;
    auto it = cache.kernels.find(signature);
    if (it != cache.kernels.end()) {
        cache.hits++;
        if (!it->second.error.empty()) { cache.last_error = it->second.error; return nullptr; }
        return &it->second;
    }
    JitKernel& k = cache.kernels[signature];
    auto t0 = std::chrono::steady_clock::now();
    std::string cubin;
    if (!jit_compile_cubin(source, cubin, k.error)) {
        cache.failures++;
        cache.last_error = k.error;
        return nullptr;
    }
    cudaError_t e = cudaLibraryLoadData(&k.lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e == cudaSuccess) e = cudaLibraryGetKernel(&k.fn, k.lib, entry);
    if (e != cudaSuccess) {
        cudaGetLastError();
        k.error = std::string("loading the compiled kernel failed: ") + cudaGetErrorString(e);
        k.fn = nullptr;
        cache.failures++;
        cache.last_error = k.error;
        return nullptr;
    }
    k.compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    cache.compiled++;
    return &k;
}

void jit_launch(pg_ctx* ctx, JitKernel* k, const char* name, int grid, int block, size_t smem, void** args) {
    if ((int)smem > k->max_dyn_smem) {
        PG_CUDA(cudaFuncSetAttribute((const void*)k->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k->max_dyn_smem = (int)smem;
    }
    flush_staged_inputs(ctx);
    pg_kernel_record rec;
    if (ctx->timing) {
        rec.name = name; rec.start = timing_event(ctx); rec.stop = timing_event(ctx);
        cudaEventRecord(rec.start, ctx->stream);
    }
    PG_CUDA(cudaLaunchKernel((const void*)k->fn, dim3(grid), dim3(block), args, smem, ctx->stream));
    if (ctx->timing) { cudaEventRecord(rec.stop, ctx->stream); ctx->records.push_back(rec); }
    ctx->launches++;
}

void jit_cache_free(pg_ctx* ctx) {
    delete ctx->jit;
    ctx->jit = nullptr;
}

}  // namespace pg
