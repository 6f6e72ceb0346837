"""Weak-supervision loss of the reference's training script (train.py:110-156), restated so that
``train.py``'s step can be driven on the new head.  For CUDA volumes with the default soft-max normalisation the
score runs on the fused kernels (ncnet_match_score_fwd/bwd); the other normalisations use PyTorch reductions."""
from __future__ import annotations

import numpy as np
import torch


def match_score(corr4d: torch.Tensor, normalization="softmax") -> torch.Tensor:
    """mean over cells of (max over A of norm_A(x) + max over B of norm_B(x)) / 2  (train.py:123-134)."""
    if normalization == "softmax" and corr4d.is_cuda and corr4d.dim() == 6 and corr4d.shape[1] == 1:
        from .functional import match_score_softmax
        return match_score_softmax(corr4d)            # fused kernels
    b = corr4d.size(0)
    f1, f2, f3, f4 = corr4d.shape[2:]
    nc_b_avec = corr4d.view(b, f1 * f2, f3, f4)
    nc_a_bvec = corr4d.view(b, f1, f2, f3 * f4).permute(0, 3, 1, 2)
    if normalization is None:
        norm = lambda x: x
    elif normalization == "softmax":
        norm = lambda x: torch.nn.functional.softmax(x, 1)
    elif normalization == "l1":
        norm = lambda x: x / (torch.sum(x, dim=1, keepdim=True) + 0.0001)
    else:
        raise ValueError(normalization)
    scores_b, _ = torch.max(norm(nc_b_avec), dim=1)
    scores_a, _ = torch.max(norm(nc_a_bvec), dim=1)
    return torch.mean(scores_a + scores_b) / 2


def weak_loss(model, batch, normalization="softmax", alpha=30):
    """Same call signature and side effect as train.py:110 (the batch's source images are rolled)."""
    b = batch["source_image"].size(0)
    score_pos = match_score(model(batch), normalization)
    batch["source_image"] = batch["source_image"][np.roll(np.arange(b), -1), :]
    score_neg = match_score(model(batch), normalization)
    return score_neg - score_pos


def weak_loss_from_features(model, feature_A, feature_B, normalization="softmax"):
    """weak_loss with the trunk factored out: the negative pass pairs target b with source b+1
    (rolling the images rolls their features).  Used by the tests."""
    score_pos = match_score(model.match_features(feature_A, feature_B), normalization)
    score_neg = match_score(model.match_features(torch.roll(feature_A, shifts=-1, dims=0), feature_B), normalization)
    return score_neg - score_pos


_WEIGHT_CACHE = {}


def weak_loss_stacked(model, source_pos, source_neg, target):
    """The same weak loss (train.py:135-139, soft-max normalisation) from ONE pass of the head: the B positive pairs
    (source_pos[b], target[b]) and the B negative pairs (source_neg[b], target[b]) are stacked into a batch of 2B and the
    fused score kernel sums them with weights -1/B and +1/B.  Same 2B pair evaluations as two calls of the model, half
    the kernel launches, twice the work per launch.  ``source_neg`` is the globally rolled source batch sliced like
    ``source_pos`` (ncnet_b200.dist.shard_pairs)."""
    return weak_loss_of_stacked(model, torch.cat([source_pos, source_neg], 0), torch.cat([target, target], 0))


def weak_loss_of_stacked(model, source2, target2):
    """``weak_loss_stacked`` for inputs that are already stacked: the first half of the batch holds the positive
    pairs, the second half the negative ones."""
    from .functional import match_score_softmax
    b = source2.shape[0] // 2
    key = (b, str(source2.device))
    w = _WEIGHT_CACHE.get(key)
    if w is None:
        w = _WEIGHT_CACHE[key] = torch.cat([torch.full((b,), -1.0 / b), torch.full((b,), 1.0 / b)]).to(source2.device)
    return match_score_softmax(model.match_features(source2, target2), w)
