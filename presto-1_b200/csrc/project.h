// project.h — PageProcessor: filter + projections over device-resident columns (include/prestogpu.h, "PageProcessor").
// Replaces OP/project/PageProcessor.java:139-330 and the expression classes PageFunctionCompiler generates, for the
// RowExpression subset of pg_expr.  Two execution forms share one validated node array:
//   * an interpreting kernel pair (project.cu): filter -> selected positions -> projections, used by pg_project_process and as the
//     general path of the fused operators;
//   * C source for the run-time specialised aggregation kernels (project_codegen): the expressions become straight-line code
//     inside the kernel that consumes them, so derived columns never exist in memory.
#pragma once
#include <string>
#include <vector>

#include "context.h"

namespace pg {

static const int PROJ_MAX_NODES = 48;
static const int PROJ_MAX_CHANNELS = 32;
static const int PROJ_MAX_OUTPUTS = 24;

struct PNode {           // device-visible
    int32_t op, type, precision, scale;
    int32_t channel, a0, a1, is_null;
    int64_t lo, hi;      // constant (VARCHAR: lo = offset into the constant bytes, hi = length)
    int32_t up0, up1;    // ADD / SUB / CAST: power of ten the operand is multiplied by to reach the result scale
    int32_t in_filter;   // part of the filter expression (evaluated for every row, before the projections)
    int32_t pad;
};

struct ProjCols { DCol c[PROJ_MAX_CHANNELS]; };

// The projected page of one input page: device columns owned here (or aliases of the input columns for pass-through projections
// when every row was selected).
struct ProjectedPage {
    int64_t n = 0;
    std::vector<DCol> cols;
    std::vector<DevBuf> bufs;
    std::vector<char> in_place;     // column k was written into the caller's buffer (project_materialize's `direct`)
};

}  // namespace pg

struct pg_project {
    pg_ctx* ctx = nullptr;
    std::vector<pg::PNode> nodes;
    int32_t filter = -1;
    std::vector<int32_t> proj;
    std::vector<uint8_t> const_bytes;
    pg::DevBuf d_nodes, d_bytes;
    int32_t max_channel = -1;
    std::vector<char> channel_used;
};

namespace pg {
// only_bits (optional): bitmap over the n input rows; rows whose bit is 0 are treated as rejected (used for the rows a fused
// kernel left to the general path).  want (optional): which projections to materialise (others stay DCol{} with values null).
void project_materialize(pg_project* p, const DCol* in, int ncols, int64_t n, const unsigned int* only_bits, const std::vector<char>* want,
                         ProjectedPage& out, const pg_col_out* direct = nullptr);
// gather rows `pos` of a device column into fresh device buffers (join.cu kernels); returns the value bytes written
void gather_device_col(pg_ctx* ctx, const DCol& src, const int64_t* d_pos, int64_t n, DevBuf& values, DevBuf& nulls, DevBuf& offsets, DCol& out);
void gather_device_col_into(pg_ctx* ctx, const DCol& src, const int64_t* d_pos, int64_t n, void* values, uint8_t* nulls);   // fixed width only

// ---- code generation for the specialised kernels
struct ProjCode {
    std::string decl;                 // per-row statements defining the variables below (uses PGX_IN* macros for column loads)
    std::vector<std::string> lo, hi;  // per node: expression text of the value; a long decimal is two's complement here (lo = low
                                      // 64 bits, hi = signed upper 64 bits, decimal_ops.cuh Pgx2c) because it only feeds sums / comparisons
    std::vector<std::string> null;    // per node: expression text of "is null" ("false" when it cannot be)
    bool can_overflow = false;
};
// Emits straight-line C for the nodes reachable from `roots`; in_val(channel) / in_null(channel) give the text that reads input
// column `channel` of the current row (empty null text = column without nulls).  Nodes of the filter come first and the text
// "PGX_FILTER_DONE" is emitted after them so that the caller can skip the rest for rejected rows.
bool project_codegen(const pg_project* p, const std::vector<int>& roots, const std::vector<std::string>& in_lo, const std::vector<std::string>& in_hi,
                     const std::vector<std::string>& in_null, ProjCode& out);
}  // namespace pg
