"""flashinfer_b200 — B200-native (sm_100a) paged-KV batch decode attention behind
FlashInfer's own Python API (reference flashinfer/__init__.py exports the same names).

Only the decode hot path is provided: the batch / single decode entry
points, the attention-state merge operators, the cascade wrappers built on both and the
page-table maintenance ops.  Everything runs in hand-written CUDA
behind the C-ABI library `libfi_b200.so` (include/fi_b200.h); there is no CPU fallback.
"""
from .cascade import (
    BatchDecodeWithSharedPrefixPagedKVCacheWrapper,
    MultiLevelCascadeAttentionWrapper,
    merge_state,
    merge_state_in_place,
    merge_states,
    variable_length_merge_states,
)
from .decode import (
    BatchDecodeWithPagedKVCacheWrapper,
    CUDAGraphBatchDecodeWithPagedKVCacheWrapper,
    fast_decode_plan,
    single_decode_with_kv_cache,
)
from .page import append_paged_kv_cache, get_batch_indices_positions, get_seq_lens

__version__ = "0.1.0"

__all__ = [
    "BatchDecodeWithPagedKVCacheWrapper",
    "CUDAGraphBatchDecodeWithPagedKVCacheWrapper",
    "fast_decode_plan",
    "single_decode_with_kv_cache",
    "merge_state",
    "merge_state_in_place",
    "merge_states",
    "variable_length_merge_states",
    "MultiLevelCascadeAttentionWrapper",
    "BatchDecodeWithSharedPrefixPagedKVCacheWrapper",
    "get_seq_lens",
    "get_batch_indices_positions",
    "append_paged_kv_cache",
]
