"""COCO detection metric for SSD-300 evaluation: per-class NMS decoding of the network output and
a NumPy implementation of the COCO bounding-box AP / AR protocol.

Reference: T/coco_metric.py:50-178 (MLPerf-derived ``decode_single`` + pycocotools
``COCOeval``). pycocotools is not available offline, so the evaluation protocol (greedy
matching per category and image at IoU .50:.05:.95, crowd regions as ignore, area ranges,
101-point interpolated precision, maxDets 1/10/100) is restated here; on the standard corner
cases (perfect detections, shifted boxes, duplicates, crowd) it gives the numbers COCOeval
defines (tests/test_ssd_data_and_metric.py).
"""
from __future__ import annotations

import json
from collections import defaultdict
from typing import Dict, Iterable, List, Sequence

import numpy as np

from .data import ssd_dataloader as sd

OVERLAP_CRITERIA = 0.5
MIN_SCORE = 0.05
DUMMY_SCORE = -1e5
METRIC_NAMES = ["AP", "AP50", "AP75", "APs", "APm", "APl", "ARmax1", "ARmax10", "ARmax100",
                "ARs", "ARm", "ARl"]


def calc_iou(target: np.ndarray, candidates: np.ndarray) -> np.ndarray:
    """IoU of one box against many (ymin, xmin, ymax, xmax)."""
    lt = np.maximum(target[None, :2], candidates[:, :2])
    rb = np.minimum(target[None, 2:], candidates[:, 2:])
    d = np.maximum(rb - lt, 0)
    inter = d[:, 0] * d[:, 1]
    a1 = (target[2] - target[0]) * (target[3] - target[1])
    a2 = (candidates[:, 2] - candidates[:, 0]) * (candidates[:, 3] - candidates[:, 1])
    return inter / np.maximum(a1 + a2 - inter, 1e-12)


def decode_single(bboxes_in, scores_in, criteria=OVERLAP_CRITERIA,
                  max_output=sd.MAX_NUM_EVAL_BOXES, max_num=200):
    """Per-class greedy NMS (T/coco_metric.py:118-164): classes 1.., scores above 0.05, the 200
    best candidates per class, suppression at IoU ``criteria``, the ``max_output`` best overall.
    Returns (boxes, labels, scores)."""
    b_out, s_out, l_out = [], [], []
    for cls in range(1, scores_in.shape[1]):
        score = scores_in[:, cls]
        mask = score > MIN_SCORE
        if not mask.any():
            continue
        boxes, score = bboxes_in[mask], score[mask]
        order = np.argsort(score)[-max_num:]
        keep = []
        while len(order):
            idx = order[-1]
            iou = calc_iou(boxes[idx], boxes[order])
            order = order[iou < criteria]
            keep.append(idx)
        b_out.append(boxes[keep])
        s_out.append(score[keep])
        l_out.extend([cls] * len(keep))
    if not s_out:
        return (np.zeros((1, 4), np.float32), np.zeros((1,), np.int32),
                np.full((1,), DUMMY_SCORE, np.float32))
    b = np.concatenate(b_out)
    s = np.concatenate(s_out)
    l = np.asarray(l_out)
    top = np.argsort(s)[-max_output:]
    return b[top], l[top], s[top]


def predictions_to_coco(examples: Iterable[dict]) -> List[list]:
    """[{pred_boxes [A,4] decoded & normalised, pred_scores [A,L], source_id, raw_shape}] ->
    rows [image_id, x, y, w, h, score, category_id] in pixels (T/coco_metric.py:58-74)."""
    rows = []
    for ex in examples:
        loc, label, prob = decode_single(np.asarray(ex["pred_boxes"]), np.asarray(ex["pred_scores"]))
        htot, wtot = ex["raw_shape"][0], ex["raw_shape"][1]
        for b, l, p in zip(loc, label, prob):
            rows.append([int(ex["source_id"]), b[1] * wtot, b[0] * htot, (b[3] - b[1]) * wtot,
                         (b[2] - b[0]) * htot, float(p), sd.CLASS_INV_MAP[int(l)]])
    return rows


def _iou_xywh(dets: np.ndarray, gts: np.ndarray, crowd: np.ndarray) -> np.ndarray:
    """IoU matrix [D, G] of xywh boxes; for crowd ground truth the union is the detection area."""
    dx0, dy0, dx1, dy1 = dets[:, 0], dets[:, 1], dets[:, 0] + dets[:, 2], dets[:, 1] + dets[:, 3]
    gx0, gy0, gx1, gy1 = gts[:, 0], gts[:, 1], gts[:, 0] + gts[:, 2], gts[:, 1] + gts[:, 3]
    iw = np.maximum(np.minimum(dx1[:, None], gx1[None]) - np.maximum(dx0[:, None], gx0[None]), 0)
    ih = np.maximum(np.minimum(dy1[:, None], gy1[None]) - np.maximum(dy0[:, None], gy0[None]), 0)
    inter = iw * ih
    da = (dets[:, 2] * dets[:, 3])[:, None]
    ga = (gts[:, 2] * gts[:, 3])[None]
    union = np.where(crowd[None], da, da + ga - inter)
    return inter / np.maximum(union, 1e-12)


_AREAS = {"all": (0, 1e10), "small": (0, 32 ** 2), "medium": (32 ** 2, 96 ** 2),
          "large": (96 ** 2, 1e10)}
_IOU_THRS = np.linspace(0.5, 0.95, 10)
_REC_THRS = np.linspace(0.0, 1.0, 101)
_MAX_DETS = (1, 10, 100)


def _evaluate_img(dets, gts, area_rng, max_det):
    """Greedy matching of one (image, category) pair, COCOeval.evaluateImg."""
    if not gts:
        # no ground truth of this category in the image (most pairs of a detector's output):
        # every detection is unmatched, ignored only if its area is out of range
        dets = sorted(dets, key=lambda d: -d["score"])[:max_det]
        d_area = np.array([d["bbox"][2] * d["bbox"][3] for d in dets])
        oor = (d_area < area_rng[0]) | (d_area > area_rng[1]) if dets else np.zeros(0, bool)
        T = len(_IOU_THRS)
        return dict(scores=np.array([d["score"] for d in dets]),
                    dtm=-np.ones((T, len(dets)), np.int64),
                    dt_ignore=np.repeat(oor[None, :], T, 0), num_gt=0)
    g_ignore = np.array([(g["iscrowd"] or g["area"] < area_rng[0] or g["area"] > area_rng[1])
                         for g in gts], bool)
    g_order = np.argsort(g_ignore, kind="mergesort")
    gts = [gts[i] for i in g_order]
    g_ignore = g_ignore[g_order]
    dets = sorted(dets, key=lambda d: -d["score"])[:max_det]
    T, D, G = len(_IOU_THRS), len(dets), len(gts)
    dtm = -np.ones((T, D), np.int64)
    gtm = -np.ones((T, G), np.int64)
    dt_ignore = np.zeros((T, D), bool)
    if D and G:
        crowd = np.array([bool(g["iscrowd"]) for g in gts])
        ious = _iou_xywh(np.array([d["bbox"] for d in dets], np.float64),
                         np.array([g["bbox"] for g in gts], np.float64), crowd)
        for ti, t in enumerate(_IOU_THRS):
            for di in range(D):
                best, m = min(t, 1 - 1e-10), -1
                for gi in range(G):
                    if gtm[ti, gi] >= 0 and not crowd[gi]:
                        continue
                    if m > -1 and not g_ignore[m] and g_ignore[gi]:
                        break
                    if ious[di, gi] < best:
                        continue
                    best, m = ious[di, gi], gi
                if m == -1:
                    continue
                dt_ignore[ti, di] = g_ignore[m]
                dtm[ti, di] = m
                gtm[ti, m] = di
    d_area = np.array([d["bbox"][2] * d["bbox"][3] for d in dets])
    out_of_range = (d_area < area_rng[0]) | (d_area > area_rng[1]) if D else np.zeros(0, bool)
    dt_ignore = dt_ignore | ((dtm < 0) & out_of_range[None, :])
    return dict(scores=np.array([d["score"] for d in dets]), dtm=dtm, dt_ignore=dt_ignore,
                num_gt=int((~g_ignore).sum()))


def coco_eval(gt_annotations: Sequence[dict], detections: Sequence[Sequence[float]],
              image_ids=None) -> Dict[str, float]:
    """COCO bbox metrics. ``gt_annotations``: COCO ``annotations`` entries (image_id,
    category_id, bbox xywh, area, iscrowd); ``detections``: rows [image_id, x, y, w, h, score,
    category_id]."""
    gts = defaultdict(list)
    dts = defaultdict(list)
    for a in gt_annotations:
        gts[(a["image_id"], a["category_id"])].append(
            dict(bbox=a["bbox"], area=a.get("area", a["bbox"][2] * a["bbox"][3]),
                 iscrowd=int(a.get("iscrowd", 0))))
    for r in detections:
        dts[(int(r[0]), int(r[6]))].append(dict(bbox=list(map(float, r[1:5])), score=float(r[5])))
    imgs = sorted(set(k[0] for k in list(gts) + list(dts))) if image_ids is None \
        else sorted(image_ids)
    cats = sorted(set(k[1] for k in gts))
    T, R = len(_IOU_THRS), len(_REC_THRS)
    precision = -np.ones((T, R, len(cats), len(_AREAS), len(_MAX_DETS)))
    recall = -np.ones((T, len(cats), len(_AREAS), len(_MAX_DETS)))
    for ci, cat in enumerate(cats):
        for ai, rng in enumerate(_AREAS.values()):
            # One greedy matching per image with the largest detection budget: detections are
            # visited in score order and a match never depends on later detections, so the
            # result for a smaller budget is the leading part of the same arrays.
            full = [_evaluate_img(dts.get((im, cat), []), gts.get((im, cat), []), rng,
                                  max(_MAX_DETS))
                    for im in imgs if (im, cat) in gts or (im, cat) in dts]
            npig = sum(r["num_gt"] for r in full)
            if npig == 0:
                continue
            for mi, md in enumerate(_MAX_DETS):
                scores = np.concatenate([r["scores"][:md] for r in full]) if full else np.zeros(0)
                order = np.argsort(-scores, kind="mergesort")
                dtm = np.concatenate([r["dtm"][:, :md] for r in full], 1)[:, order]
                dig = np.concatenate([r["dt_ignore"][:, :md] for r in full], 1)[:, order]
                tps = np.cumsum((dtm >= 0) & ~dig, 1).astype(np.float64)
                fps = np.cumsum((dtm < 0) & ~dig, 1).astype(np.float64)
                for ti in range(T):
                    tp, fp = tps[ti], fps[ti]
                    nd = len(tp)
                    rc = tp / npig
                    pr = tp / np.maximum(tp + fp, np.spacing(1))
                    recall[ti, ci, ai, mi] = rc[-1] if nd else 0
                    pr = pr.tolist()
                    for i in range(nd - 1, 0, -1):
                        if pr[i] > pr[i - 1]:
                            pr[i - 1] = pr[i]
                    inds = np.searchsorted(rc, _REC_THRS, side="left")
                    q = np.zeros(R)
                    for ri, pi in enumerate(inds):
                        if pi < nd:
                            q[ri] = pr[pi]
                    precision[ti, :, ci, ai, mi] = q

    def _ap(iou=None, area="all", md=100):
        ai, mi = list(_AREAS).index(area), _MAX_DETS.index(md)
        p = precision[:, :, :, ai, mi]
        if iou is not None:
            p = p[[int(round((iou - 0.5) / 0.05))]]
        p = p[p > -1]
        return float(p.mean()) if p.size else -1.0

    def _ar(area="all", md=100):
        ai, mi = list(_AREAS).index(area), _MAX_DETS.index(md)
        r = recall[:, :, ai, mi]
        r = r[r > -1]
        return float(r.mean()) if r.size else -1.0

    stats = [_ap(), _ap(0.5), _ap(0.75), _ap(area="small"), _ap(area="medium"), _ap(area="large"),
             _ar(md=1), _ar(md=10), _ar(md=100), _ar("small"), _ar("medium"), _ar("large")]
    return {"COCO/" + k: v for k, v in zip(METRIC_NAMES, stats)}


def compute_map(labels_and_predictions: Iterable[dict], val_json_file: str) -> Dict[str, float]:
    """T/coco_metric.py:50-100: decode + NMS every example, then COCO AP against the
    annotation file (``instances_val2017.json`` layout)."""
    with open(val_json_file) as f:
        ann = json.load(f)
    rows = predictions_to_coco(list(labels_and_predictions))
    stats = coco_eval(ann["annotations"], rows)
    print("Current AP: {:.5f}".format(stats["COCO/AP"]))
    return stats
