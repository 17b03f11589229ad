"""Page-table helpers used by the decode path (reference flashinfer/page.py:224-247)."""
from __future__ import annotations

import torch


def get_seq_lens(
    kv_indptr: torch.Tensor, kv_last_page_len: torch.Tensor, page_size: int
) -> torch.Tensor:
    r"""Convert KV indptr and last page length to sequence lengths.

    ``kv_len[i] = max(indptr[i+1] - indptr[i] - 1, 0) * page_size + last_page_len[i]``
    (reference flashinfer/page.py:243-246).
    """
    return (
        torch.clamp(kv_indptr[1:] - kv_indptr[:-1] - 1, min=0) * page_size
        + kv_last_page_len
    )
