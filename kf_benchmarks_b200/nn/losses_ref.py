"""PyTorch restatements of the non-classification loss heads (oracles of csrc/loss_kernels.cu
and the ``--device=cpu`` path). ``compute`` returns (loss, metric1, metric2)."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def ssd_gather(graph, vals):
    """[N, 8732, 4] locations and [N, 8732, L] class scores from the head convolutions, in the
    anchor order of T/ssd_dataloader.py:35-79 (map, default box, row, column)."""
    spec = graph.loss
    L = spec["num_classes"]
    locs, confs = [], []
    for loc_id, conf_id, d in zip(spec["loc"], spec["conf"], spec["num_dboxes"]):
        lt, ct = vals[loc_id].float(), vals[conf_id].float()
        n, s = lt.shape[0], lt.shape[1]
        locs.append(lt[..., :d * 4].reshape(n, s, s, d, 4).permute(0, 3, 1, 2, 4)
                    .reshape(n, d * s * s, 4))
        confs.append(ct[..., :d * L].reshape(n, s, s, d, L).permute(0, 3, 1, 2, 4)
                     .reshape(n, d * s * s, L))
    return torch.cat(locs, 1), torch.cat(confs, 1)


def ssd_loss(pred_loc, pred_label, gt_loc, gt_label, num_matched, negs_per_positive=3):
    """T/models/ssd_model.py:312-384."""
    gt = gt_label.reshape(gt_label.shape[0], -1).long()
    mask = (gt > 0).float()
    huber = F.smooth_l1_loss(pred_loc, gt_loc, reduction="none", beta=1.0).sum(2)
    box = (huber * mask).sum(1) / num_matched
    n, a, L = pred_label.shape
    xent = F.cross_entropy(pred_label.reshape(n * a, L), gt.clamp(0, L - 1).reshape(-1),
                           reduction="none").reshape(n, a)
    neg = xent.detach() * (1 - mask)
    order = torch.argsort(neg, dim=1, descending=True, stable=True)
    rank = torch.argsort(order, dim=1, stable=True)
    k = (num_matched.to(torch.int32) * negs_per_positive).clamp(max=a)
    topk = (rank < k[:, None]).float()
    cls = (xent * (mask + topk)).sum(1) / num_matched
    return cls.mean() + box.mean(), cls.mean(), box.mean()


def compute(kind, graph, vals, labels):
    if kind == "ssd":
        loc, conf = ssd_gather(graph, vals)
        gt_loc, gt_label, nm = labels
        total, _, _ = ssd_loss(loc, conf, gt_loc.float(), gt_label.float(), nm.float(),
                               graph.loss.get("negs_per_positive", 3))
        return total, 0, 0
    if kind == "ctc":
        from . import ctc_ref
        return ctc_ref.loss(graph, vals, labels)
    raise NotImplementedError("loss head %r" % kind)
