"""presto-1_b200 — B200-native hash-aggregation / hash-join operator path behind the reference's operator API.

The package directory name contains a hyphen (it follows the reference repo's name); import it with
`importlib.import_module("presto-1_b200")` or through the `presto_b200` shim at the repository root.
"""
from . import _native as native
from ._native import (AGG_AVG, AGG_COUNT, AGG_COUNT_COL, AGG_MAX, AGG_MIN, AGG_SUM, BIGINT, BOOLEAN, DATE, DOUBLE, INTEGER,
                      LONG_DECIMAL, SHORT_DECIMAL, VARBINARY, VARCHAR, PrestoGpuError)
from .blocks import Block, OutBlock, Page
from .context import GpuContext, default_context
from .groupby import GroupByHash, GroupByIdBlock
from .accumulators import GroupedAccumulator, final_type, intermediate_types
from .join import LookupSource, PagesIndex
from .operators import (FINAL, INTERMEDIATE, PARTIAL, SINGLE, AggregatorFactory, HashAggregationOperator, HashBuilderOperator,
                        DistinctLimitOperator, HashSemiJoinOperator, LookupJoinOperator, LookupOuterOperator, MarkDistinctOperator, SetBuilderOperator, take_positions)
JOIN_INNER, JOIN_PROBE_OUTER = native.JOIN_INNER, native.JOIN_PROBE_OUTER
from .exchange import DevicePartitioner, HashPartitionExchange
from .plans import PartialFinalAggregation, PartitionedAggregation, PartitionedHashJoin, final_aggregators
from . import pages_serde
from .expressions import Expr, PageProcessor, ScanFilterAndProjectOperator
