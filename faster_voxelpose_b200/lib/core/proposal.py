"""Drop-in for reference ``lib/core/proposal.py``: ``nms2D(prob_map, max_num)`` (reference :54-76)
as one CUDA kernel (3x3 max-equality NMS + top-k + index conversion)."""
from __future__ import annotations

from ... import ops


def nms2D(prob_map, max_num):
    """prob_map [B,1,X,Y] -> (topk_values [B,P], topk_index [B,P,2], topk_flatten_index [B,P]).

    Ties are ordered larger value first, then smaller flat index (torch.topk leaves tie order
    unspecified).  The 2-D index uses the reference's divisor (F8)."""
    return ops.nms_topk(prob_map, max_num)
