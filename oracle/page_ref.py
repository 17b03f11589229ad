"""CPU restatement of the reference's page-table maintenance ops (test infrastructure only: imported
by tests/ and nothing else; the product path never touches it).

* get_batch_indices_positions — flashinfer/page.py:169-221 with the Triton kernel of
  flashinfer/triton/page.py:21-42: for request i, entries j in [indptr[i], indptr[i+1]) get
  batch_indices[j] = i and positions[j] = j + seq_lens[i] - indptr[i+1].
* append_paged_kv_cache — AppendPagedKVCacheKernel, include/flashinfer/page.cuh:254-287:
  page_iter, entry = divmod(kv_indptr[b] * page_size + pos, page_size); the (token, head) row is
  copied to page kv_indices[page_iter], slot `entry`, for K and for V.

Pinned by the known answers printed in the reference's own docstrings
(flashinfer/page.py:190-199 and :365-383), see tests/test_page.py.
"""
from __future__ import annotations

import numpy as np
import torch


def get_batch_indices_positions_ref(append_indptr, seq_lens, nnz: int):
    indptr = np.asarray(append_indptr, dtype=np.int64)
    lens = np.asarray(seq_lens, dtype=np.int64)
    batch_indices = np.zeros(nnz, dtype=np.int32)
    positions = np.zeros(nnz, dtype=np.int32)
    for i in range(len(indptr) - 1):
        for j in range(indptr[i], min(indptr[i + 1], nnz)):
            batch_indices[j] = i
            positions[j] = j + lens[i] - indptr[i + 1]
    return batch_indices, positions


def append_paged_kv_cache_ref(append_key, append_value, batch_indices, positions, paged_kv_cache,
                              kv_indices, kv_indptr, kv_last_page_len, kv_layout="NHD"):
    """In place on CPU tensors (any dtype: rows are copied, never converted)."""
    if isinstance(paged_kv_cache, tuple):
        k_cache, v_cache = paged_kv_cache
    else:
        k_cache, v_cache = paged_kv_cache[:, 0], paged_kv_cache[:, 1]
    page_size = k_cache.shape[1] if kv_layout == "NHD" else k_cache.shape[2]
    bi = [int(x) for x in batch_indices]
    ps = [int(x) for x in positions]
    ind = [int(x) for x in kv_indices]
    ptr = [int(x) for x in kv_indptr]
    kb, vb = k_cache.view(torch.uint8) if k_cache.element_size() == 1 else k_cache, \
        v_cache.view(torch.uint8) if v_cache.element_size() == 1 else v_cache
    ak = append_key.view(torch.uint8) if append_key.element_size() == 1 else append_key
    av = append_value.view(torch.uint8) if append_value.element_size() == 1 else append_value
    for j in range(len(bi)):
        page_iter, entry = divmod(ptr[bi[j]] * page_size + ps[j], page_size)
        page = ind[page_iter]
        if kv_layout == "NHD":
            kb[page, entry] = ak[j]
            vb[page, entry] = av[j]
        else:
            kb[page, :, entry] = ak[j]
            vb[page, :, entry] = av[j]
