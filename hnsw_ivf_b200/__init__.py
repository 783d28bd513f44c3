"""hnsw_ivf_b200 -- B200-native batched IVF-Flat search (one hot path of Quaternijkon/hnsw-ivf).

Python host mirror of the reference's interface for this path. The product is the C-ABI shared
library ``lib/libb200ivf.so`` (hand-written sm_100a CUDA, see csrc/); this module only binds it
with ctypes, the way the reference's Python layer binds its C++ (faiss/python/class_wrappers.py:300
``replacement_search``: numpy in, (D, I) out). There is NO CPU fallback: importing works without a
GPU (so the symbols can be checked), but every compute call raises if the library or a CUDA device
is missing.
"""
from .binding import (  # noqa: F401
    METRIC_INNER_PRODUCT,
    METRIC_L2,
    B200Error,
    IDSelector,
    IDSelectorAnd,
    IDSelectorArray,
    IDSelectorBatch,
    IDSelectorBitmap,
    IDSelectorNot,
    IDSelectorOr,
    IDSelectorRange,
    IDSelectorXOr,
    IndexIVFFlat,
    IndexIVFStats,
    QueryLatencyStats,
    SearchParametersIVF,
    build_library,
    kernel_launches,
    tensor_core_launches,
    library_path,
    load_library,
    merge_knn_results,
    read_index,
    write_index,
)
from .sharded import (IndexShardsIVFFlat, ShardedSearchModel, list_range_of_rank,  # noqa: F401
                      nccl_unique_id)

__all__ = [
    "METRIC_INNER_PRODUCT", "METRIC_L2", "B200Error", "IndexIVFFlat", "IndexIVFStats",
    "IDSelector", "IDSelectorRange", "IDSelectorBatch", "IDSelectorArray", "IDSelectorBitmap",
    "IDSelectorNot", "IDSelectorAnd", "IDSelectorOr", "IDSelectorXOr", "SearchParametersIVF",
    "QueryLatencyStats",
    "IndexShardsIVFFlat", "build_library", "kernel_launches", "tensor_core_launches", "library_path", "load_library", "merge_knn_results", "read_index", "write_index",
    "list_range_of_rank", "ShardedSearchModel", "nccl_unique_id",
]
