"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/prestogpu.h declares,
and refuses to work without a CUDA device (there is no CPU fallback in the product path)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import presto_b200 as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    lib = P.native.lib()
    declared = P.native.declared_symbols()
    assert len(declared) >= 45
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    # and the ctypes prototypes cover the header
    assert set(declared) <= set(P.native._PROTOS) | {"pg_version"}, sorted(set(declared) - set(P.native._PROTOS))


def test_header_cites_reference_interfaces():
    text = open(os.path.join(ROOT, "include", "prestogpu.h")).read()
    for ref in ("OP/GroupByHash.java", "OP/BigintGroupByHash.java", "OP/MultiChannelGroupByHash.java", "GroupedAccumulator.java",
                "OP/HashAggregationOperator.java", "OP/HashBuilderOperator.java", "OP/PagesHash.java", "OP/LookupSource.java",
                "OP/JoinProbe.java", "LocalPartitionGenerator.java"):
        assert ref in text, ref
    assert re.search(r'extern "C"', text)
    assert "#include <torch" not in text and "at::Tensor" not in text    # plain pointers and sizes only


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(P.PrestoGpuError) as e:
        P.GpuContext(0)
    assert e.value.code == P.native.PG_E_CUDA


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "presto-1_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in src and "liboracle" not in src and "oracle/" not in src.replace("oracle/oracle_groupby.h type_hash", ""), f


def test_block_layouts_match_the_reference_blocks():
    b = P.Block.int128_array([5, -5, None, 2**100])
    assert b.type == P.LONG_DECIMAL and b.n == 4 and b.to_list() == [5, -5, None, 2**100]
    assert b.values[2] == 5 and b.values[3] == np.uint64(1 << 63)            # sign-magnitude: sign bit in the high word
    v = P.Block.variable_width(["ab", None, "", "xyz"])
    assert v.offsets.tolist() == [0, 2, 2, 2, 5] and v.nulls.tolist() == [0, 1, 0, 0] and v.to_list() == [b"ab", None, b"", b"xyz"]
    d = P.Block.double_array([1.5, -0.0])
    assert d.type == P.DOUBLE and d.values.dtype == np.int64 and d.to_list() == [1.5, -0.0]
    c = d.pg_col()
    assert c.mem == P.native.MEM_HOST and c.n == 2 and c.values == d.values.ctypes.data
