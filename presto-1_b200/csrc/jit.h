// jit.h — run-time kernel specialisation: NVRTC (loaded lazily with dlopen, so the library itself links against nothing
// but the static CUDA runtime) compiles a plan-specific source to an sm_100a cubin, the CUDA runtime's library API loads
// it, and the kernel handle is cached per context under the plan's signature.  The reference does the same thing on the
// JVM: one generated accumulator / hash-strategy class per plan (AccumulatorCompiler.java:261-330, JoinCompiler.java:190-238).
#pragma once
#include <map>
#include <string>

#include "context.h"

namespace pg {

struct JitKernel {
    cudaLibrary_t lib = nullptr;
    cudaKernel_t fn = nullptr;
    std::string error;        // non-empty: compilation failed (cached so that it is not retried for every page)
    double compile_ms = 0;
    int max_dyn_smem = 0;     // value last set with cudaFuncSetAttribute
};

struct JitCache {
    std::map<std::string, JitKernel> kernels;
    int64_t compiled = 0, hits = 0, failures = 0;
    int lowcard_groups_hint = 0;     // groups the last discovering few-groups kernel found (sizes the next one's private tables)
    std::string last_error;
    ~JitCache();
};

// true when PRESTOGPU_JIT is unset or not "0" and NVRTC can be loaded
bool jit_available(std::string* why_not = nullptr);

// Returns the cached or freshly compiled kernel `entry` of `source` (nullptr on failure; the reason is in cache.last_error).
JitKernel* jit_get(pg_ctx* ctx, const std::string& signature, const std::string& source, const char* entry);

// Launch with launch counting and optional per-kernel timing, like PG_LAUNCH.
void jit_launch(pg_ctx* ctx, JitKernel* k, const char* name, int grid, int block, size_t smem, void** args);

}  // namespace pg
