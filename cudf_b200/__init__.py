"""cudf_b200: B200-native hash groupby-aggregate, drop-in for libcudf's cudf::groupby::groupby::aggregate hash path.

Python surface mirrors pylibcudf (GroupBy, GroupByRequest, aggregation factories, Column, Table); every call goes through
the C ABI in include/bgb.h into hand-written sm_100a kernels (cudf_b200/csrc). No CPU fallback exists.
"""
from . import aggregation  # noqa: F401
from ._lib import BgbError, CudaError, LogicError, UnsupportedError  # noqa: F401
from .aggregation import NullPolicy  # noqa: F401
from .column import Column, Table, TypeId  # noqa: F401
from .groupby import GroupBy, GroupByRequest, HostColumnView, can_use_hash_groupby, groupby_host  # noqa: F401
from . import partitioning  # noqa: F401
from .streaming import StreamingAggregationRequest, StreamingGroupBy, is_streaming_groupby_supported  # noqa: F401
