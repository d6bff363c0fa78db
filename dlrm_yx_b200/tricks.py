"""The reference's two compositional embedding variants on top of the libdlrmx lookup:

* `QREmbeddingBag`  -- tricks/qr_embedding_bag.py: two bags per lookup, one indexed by the quotient and one by the
                       remainder of the id, combined by mult / add / concat (dlrm_s_pytorch.py:281-289).
* `PrEmbeddingBag`  -- tricks/md_embedding_bag.py:52-70: a narrow table followed by a bias-free projection to the base
                       dimension (mixed-dimension embeddings, dlrm_s_pytorch.py:290-299).

Both are thin host-side compositions: every lookup, its backward and its sparse update run in the same kernels as
`EmbeddingBagSum` (embedding.py); the element-wise combine / the projection GEMM are ordinary torch ops, as in the
reference.  mode="sum" only (what DLRM_Net asks for); no CPU path.
"""
import numpy as np
import torch
from torch import nn

from .embedding import EmbeddingBagSum


class QREmbeddingBag(nn.Module):
    """QREmbeddingBag(num_categories, embedding_dim, num_collisions, operation="mult", mode="sum", sparse=True).
    `weight_q` [ceil(n / c), d0] and `weight_r` [c, d1] are the reference's parameter names."""

    def __init__(self, num_categories, embedding_dim, num_collisions, operation="mult", max_norm=None, norm_type=2.,
                 scale_grad_by_freq=False, mode="sum", sparse=True, _weight=None, device=None):
        super().__init__()
        assert operation in ["concat", "mult", "add"], "Not valid operation!"
        if mode != "sum" or max_norm is not None or scale_grad_by_freq:
            raise ValueError("QREmbeddingBag here: mode='sum', no max_norm, no scale_grad_by_freq (DLRM_Net's use)")
        self.num_categories = num_categories
        self.embedding_dim = ([embedding_dim, embedding_dim] if isinstance(embedding_dim, int) or len(embedding_dim) == 1
                              else list(embedding_dim))
        self.num_collisions, self.operation, self.mode, self.sparse = num_collisions, operation, mode, sparse
        if operation in ("add", "mult"):
            assert self.embedding_dim[0] == self.embedding_dim[1], "Embedding dimensions do not match!"
        self.num_embeddings = [int(np.ceil(num_categories / num_collisions)), num_collisions]
        if _weight is None:
            # the reference's reset_parameters: nn.init.uniform_(w, sqrt(1/n)) i.e. U(sqrt(1/n), 1)
            a = float(np.sqrt(1 / num_categories))
            _weight = [torch.empty((self.num_embeddings[k], self.embedding_dim[k]), device=device).uniform_(a, 1.0)
                       for k in range(2)]
        self.q = EmbeddingBagSum(self.num_embeddings[0], self.embedding_dim[0], _weight=_weight[0])
        self.r = EmbeddingBagSum(self.num_embeddings[1], self.embedding_dim[1], _weight=_weight[1])

    @property
    def weight_q(self):
        return self.q.weight

    @property
    def weight_r(self):
        return self.r.weight

    def forward(self, input, offsets=None, per_sample_weights=None):
        input_q = torch.div(input, self.num_collisions, rounding_mode="floor")     # (input / c).long() for ids >= 0
        input_r = torch.remainder(input, self.num_collisions)
        embed_q = self.q(input_q, offsets, per_sample_weights)
        embed_r = self.r(input_r, offsets, per_sample_weights)
        if self.operation == "concat":
            return torch.cat((embed_q, embed_r), dim=1)
        if self.operation == "add":
            return embed_q + embed_r
        return embed_q * embed_r


class PrEmbeddingBag(nn.Module):
    """PrEmbeddingBag(num_embeddings, embedding_dim, base_dim): `.embs` (the table) and `.proj` as in the reference."""

    def __init__(self, num_embeddings, embedding_dim, base_dim, device=None):
        super().__init__()
        w = torch.empty((num_embeddings, embedding_dim), device=device)
        torch.nn.init.xavier_uniform_(w)
        self.embs = EmbeddingBagSum(num_embeddings, embedding_dim, _weight=w)
        if embedding_dim < base_dim:
            self.proj = nn.Linear(embedding_dim, base_dim, bias=False, device=device)
            torch.nn.init.xavier_uniform_(self.proj.weight)
        elif embedding_dim == base_dim:
            self.proj = nn.Identity()
        else:
            raise ValueError("Embedding dim " + str(embedding_dim) + " > base dim " + str(base_dim))

    def forward(self, input, offsets=None, per_sample_weights=None):
        return self.proj(self.embs(input, offsets, per_sample_weights))
