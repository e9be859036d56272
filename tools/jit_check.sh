#!/bin/bash
# compile the run-time specialised wide-aggregation body with nvcc for a representative plan (syntax / register check without a GPU)
set -e
mkdir -p /tmp/jc
D=$(cd "$(dirname "$0")/../presto-1_b200/csrc/jit" && pwd)
for BIDX in 0 1; do for RANGE in 0 1; do for QUOTA in 0 1; do
cat > /tmp/jc/w.cu <<EOS
#define PGW_BUCKET_WORDS 8
#define PGW_HASKEYNULL 0
#define PGW_FLAGWORD 2
#define PGW_VEC 4
#define PGW_RANGE $RANGE
#define PGW_BIDX $BIDX
#define PGW_AGGS(X) X(0, CNT, NONE, 0, 0, 0, C64, 3, 4, 0) X(1, SUMDEC, I64, 0, 0, 0, FLAG, 0, 4, 0) X(2, SUMF64, F64, 0, 0, 0, FLAG, 0, 6, 1)
#define PGW_HASONLY 0
#define PGW_QUOTA $QUOTA
#include "$D/wide_args.h"
#include "$D/wide_body.cuh"
EOS
nvcc -arch=sm_100a -std=c++17 -cubin -Xptxas -v -o /tmp/jc/w.cubin /tmp/jc/w.cu 2>&1 | grep -E "error|registers|spill" || true
done; done; done
