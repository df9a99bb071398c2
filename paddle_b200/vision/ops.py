"""paddle.vision.ops (reference: python/paddle/vision/ops.py; kernels under paddle/phi/kernels/{cpu,gpu}/ for nms,
roi_align, roi_pool, psroi_pool, deformable_conv, box_coder, prior_box, yolo_box, yolo_loss, matrix_nms,
distribute_fpn_proposals, generate_proposals).

Detection operators written as vectorised tensor programs, so they run on whatever device the inputs live on and are
differentiable wherever the reference's are (roi_align, roi_pool, psroi_pool, deform_conv2d, yolo_loss).  Box layout is
``[x1, y1, x2, y2]`` throughout.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as TF

from ..nn import initializer as I
from ..nn.layer.layers import Layer

__all__ = ["yolo_loss", "yolo_box", "prior_box", "box_coder", "deform_conv2d", "DeformConv2D", "distribute_fpn_proposals",
           "generate_proposals", "read_file", "decode_jpeg", "roi_pool", "RoIPool", "psroi_pool", "PSRoIPool", "roi_align",
           "RoIAlign", "nms", "matrix_nms"]


# ------------------------------------------------------------------------------------------------------------- helpers
def _pair(v):
    return (int(v), int(v)) if isinstance(v, int) else (int(v[0]), int(v[1]))


def _cround(v):
    """C's round(): halves go away from zero (Python's round and torch.round go to even)."""
    return int(math.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)


def _box_area(boxes, offset=0.0):
    return (boxes[..., 2] - boxes[..., 0] + offset).clamp_min(0) * (boxes[..., 3] - boxes[..., 1] + offset).clamp_min(0)


def _iou_matrix(a, b, offset=0.0):
    """Pairwise IoU of [n, 4] and [m, 4]; ``offset`` = 1 for integer pixel boxes (un-normalised)."""
    lt = torch.maximum(a[:, None, :2], b[None, :, :2])
    rb = torch.minimum(a[:, None, 2:], b[None, :, 2:])
    wh = (rb - lt + offset).clamp_min(0)
    inter = wh[..., 0] * wh[..., 1]
    union = _box_area(a, offset)[:, None] + _box_area(b, offset)[None, :] - inter
    return torch.where(union > 0, inter / union.clamp_min(1e-30), torch.zeros_like(inter))


def _greedy_nms(boxes, iou_threshold, offset=0.0, eta=1.0):
    """Indices kept by greedy suppression in the given order.  ``eta`` < 1 shrinks the threshold after every kept box
    while it is above 0.5 (adaptive NMS of the proposal generator)."""
    n = boxes.shape[0]
    if n == 0:
        return torch.zeros(0, dtype=torch.int64, device=boxes.device)
    iou = _iou_matrix(boxes, boxes, offset).cpu().numpy()
    suppressed = np.zeros(n, dtype=bool)
    keep, thr = [], float(iou_threshold)
    for i in range(n):
        if suppressed[i]:
            continue
        keep.append(i)
        suppressed |= iou[i] > thr
        suppressed[i] = True
        if eta < 1.0 and thr > 0.5:
            thr *= eta
    return torch.as_tensor(keep, dtype=torch.int64, device=boxes.device)


def _split_by_counts(boxes, boxes_num, batch):
    """Image index of every box from the per-image counts."""
    if boxes_num is None:
        if batch != 1:
            raise ValueError("boxes_num is required when the batch holds more than one image")
        return torch.zeros(boxes.shape[0], dtype=torch.int64, device=boxes.device)
    counts = boxes_num.to(torch.int64).reshape(-1)
    if int(counts.sum()) != boxes.shape[0]:
        raise ValueError(f"boxes_num sums to {int(counts.sum())} but there are {boxes.shape[0]} boxes")
    return torch.repeat_interleave(torch.arange(counts.numel(), device=boxes.device), counts.to(boxes.device))


# ----------------------------------------------------------------------------------------------------------------- NMS
def nms(boxes, iou_threshold=0.3, scores=None, category_idxs=None, categories=None, top_k=None):
    """Indices of the boxes kept by non-maximum suppression.  Without ``scores`` the boxes are taken in the given
    order; with ``scores`` in falling score order; with ``category_idxs`` + ``categories`` suppression is done inside
    each category and the survivors of all categories come back sorted by score (cut to ``top_k``)."""
    if scores is None:
        return _greedy_nms(boxes, iou_threshold)
    if category_idxs is None:
        order = torch.argsort(scores, descending=True, stable=True)
        return order[_greedy_nms(boxes[order], iou_threshold)]
    if top_k is not None:
        assert top_k <= scores.shape[0], "top_k should be smaller equal than the number of boxes"
    assert categories is not None, "if category_idxs is given, categories which is a list of unique id of all categories is necessary"
    mask = torch.zeros_like(scores, dtype=torch.bool)
    for category_id in categories:
        idxs = torch.nonzero(category_idxs == category_id).reshape(-1)
        if idxs.numel() == 0:
            continue
        order = torch.argsort(scores[idxs], descending=True, stable=True)
        kept = idxs[order][_greedy_nms(boxes[idxs][order], iou_threshold)]
        mask[kept] = True
    keep = torch.nonzero(mask).reshape(-1)
    keep = keep[torch.argsort(scores[keep], descending=True, stable=True)]
    return keep[:top_k] if top_k is not None else keep


def matrix_nms(bboxes, scores, score_threshold, post_threshold, nms_top_k, keep_top_k, use_gaussian=False,
               gaussian_sigma=2.0, background_label=0, normalized=True, return_index=False, return_rois_num=True, name=None):
    """Matrix NMS (SOLOv2): instead of removing boxes, every score is decayed by its overlap with higher-scoring boxes
    of the same class: decay_j = min_i f(iou_ij) / f(max overlap of i), f linear (1 - iou) or gaussian.

    ``bboxes`` [N, M, 4], ``scores`` [N, C, M].  Returns ``out`` [K, 6] rows (label, score, x1, y1, x2, y2) sorted by
    score inside each image, then (optionally) ``index`` [K, 1] into the flattened N*M boxes and ``rois_num`` [N]."""
    n, c, m = scores.shape
    offset = 0.0 if normalized else 1.0
    outs, indices, counts = [], [], []
    for b in range(n):
        dets = []
        for cls in range(c):
            if cls == background_label:
                continue
            s = scores[b, cls]
            cand = torch.nonzero(s > score_threshold).reshape(-1)
            if cand.numel() == 0:
                continue
            order = torch.argsort(s[cand], descending=True, stable=True)
            if nms_top_k > -1:
                order = order[:nms_top_k]
            cand = cand[order]
            cs, cb = s[cand], bboxes[b, cand]
            iou = torch.triu(_iou_matrix(cb, cb, offset), diagonal=1)             # iou[i, j], i higher-scoring than j
            max_iou = iou.max(0).values                                            # largest overlap each box suffers
            if use_gaussian:
                decay = torch.exp((max_iou[:, None] ** 2 - iou ** 2) * gaussian_sigma)
            else:
                decay = (1 - iou) / (1 - max_iou[:, None])
            decay = torch.where(torch.triu(torch.ones_like(iou), diagonal=1) > 0, decay, torch.ones_like(decay)).min(0).values
            ds = cs * decay
            ok = ds > post_threshold
            if ok.any():
                lab = torch.full((int(ok.sum()), 1), float(cls), dtype=bboxes.dtype, device=bboxes.device)
                dets.append((torch.cat([lab, ds[ok, None], cb[ok]], 1), cand[ok] + b * m))
        if dets:
            det = torch.cat([d[0] for d in dets])
            idx = torch.cat([d[1] for d in dets])
            order = torch.argsort(det[:, 1], descending=True, stable=True)
            if keep_top_k > -1:
                order = order[:keep_top_k]
            outs.append(det[order])
            indices.append(idx[order])
            counts.append(order.numel())
        else:
            counts.append(0)
    out = torch.cat(outs) if outs else torch.zeros((0, 6), dtype=bboxes.dtype, device=bboxes.device)
    index = (torch.cat(indices) if indices else torch.zeros(0, dtype=torch.int64, device=bboxes.device)).reshape(-1, 1)
    rois_num = torch.as_tensor(counts, dtype=torch.int32, device=bboxes.device)
    result = [out]
    if return_index:
        result.append(index)
    if return_rois_num:
        result.append(rois_num)
    return tuple(result) if len(result) > 1 else out


# --------------------------------------------------------------------------------------------------------------- boxes
def box_coder(prior_box, prior_box_var, target_box, code_type="encode_center_size", box_normalized=True, axis=0, name=None):
    """Encode boxes against priors as (dx, dy, log dw, log dh) / variance, or decode such offsets back to boxes.

    encode: ``target_box`` [N, 4], ``prior_box`` [M, 4] -> [N, M, 4].  decode: ``target_box`` [N, M, 4]; the priors
    run along dim 1 when ``axis=0`` (``prior_box`` [M, 4]) and along dim 0 when ``axis=1`` (``prior_box`` [N, 4]).
    ``prior_box_var``: None, 4 numbers, or a tensor shaped like ``prior_box``."""
    norm = 0.0 if box_normalized else 1.0
    pw = prior_box[:, 2] - prior_box[:, 0] + norm
    ph = prior_box[:, 3] - prior_box[:, 1] + norm
    pcx = prior_box[:, 0] + pw / 2
    pcy = prior_box[:, 1] + ph / 2
    if prior_box_var is None:
        var = None
    elif isinstance(prior_box_var, torch.Tensor):
        var = prior_box_var.to(target_box.dtype)
    else:
        if len(prior_box_var) != 4:
            raise ValueError("prior_box_var should hold 4 values")
        var = torch.as_tensor(list(prior_box_var), dtype=target_box.dtype, device=target_box.device)
    if code_type == "encode_center_size":
        tw = target_box[:, 2] - target_box[:, 0] + norm
        th = target_box[:, 3] - target_box[:, 1] + norm
        tcx = (target_box[:, 0] + target_box[:, 2]) / 2
        tcy = (target_box[:, 1] + target_box[:, 3]) / 2
        out = torch.stack([(tcx[:, None] - pcx[None]) / pw[None], (tcy[:, None] - pcy[None]) / ph[None],
                           torch.log(torch.abs(tw[:, None] / pw[None])), torch.log(torch.abs(th[:, None] / ph[None]))], -1)
        if var is not None:
            out = out / (var[None] if var.dim() == 2 else var)
        return out
    if code_type != "decode_center_size":
        raise ValueError("code_type should be encode_center_size or decode_center_size")
    if target_box.dim() != 3:
        raise ValueError("decode_center_size expects a 3-D target_box")
    expand = (lambda t: t[None, :]) if axis == 0 else (lambda t: t[:, None])
    pw_, ph_, pcx_, pcy_ = expand(pw), expand(ph), expand(pcx), expand(pcy)
    t = target_box
    if var is not None:
        t = t * (expand(var) if var.dim() == 2 else var)
    cx = t[..., 0] * pw_ + pcx_
    cy = t[..., 1] * ph_ + pcy_
    w = torch.exp(t[..., 2]) * pw_
    h = torch.exp(t[..., 3]) * ph_
    return torch.stack([cx - w / 2, cy - h / 2, cx + w / 2 - norm, cy + h / 2 - norm], -1)


def prior_box(input, image, min_sizes, max_sizes=None, aspect_ratios=[1.0], variance=[0.1, 0.1, 0.2, 0.2], flip=False,  # noqa: B006
              clip=False, steps=[0.0, 0.0], offset=0.5, min_max_aspect_ratios_order=False, name=None):  # noqa: B006
    """SSD prior boxes for one feature map: (boxes [H, W, P, 4] normalised to the image, variances [H, W, P, 4]).
    Per min size: one box per aspect ratio (1 first, then the given ones and - with ``flip`` - their reciprocals),
    plus a sqrt(min * max) square when ``max_sizes`` is given; ``min_max_aspect_ratios_order`` puts the two squares
    first (Caffe order)."""
    min_sizes = [float(v) for v in (min_sizes if isinstance(min_sizes, (list, tuple)) else [min_sizes])]
    max_sizes = [float(v) for v in (max_sizes or [])]
    if max_sizes and len(max_sizes) != len(min_sizes):
        raise ValueError("max_sizes and min_sizes should have same length")
    ratios = [1.0]
    for ar in (aspect_ratios if isinstance(aspect_ratios, (list, tuple)) else [aspect_ratios]):
        ar = float(ar)
        if all(abs(ar - r) >= 1e-6 for r in ratios):
            ratios.append(ar)
            if flip:
                ratios.append(1.0 / ar)
    fh, fw = input.shape[2], input.shape[3]
    ih, iw = image.shape[2], image.shape[3]
    steps = list(steps) if isinstance(steps, (list, tuple)) else [steps, steps]
    step_w, step_h = (iw / fw, ih / fh) if steps[0] == 0 or steps[1] == 0 else (float(steps[0]), float(steps[1]))
    sizes = []                                                          # (half width, half height) per prior
    for s, mn in enumerate(min_sizes):
        squares = [(mn / 2, mn / 2)]
        if max_sizes:
            half = math.sqrt(mn * max_sizes[s]) / 2
            squares.append((half, half))
        if min_max_aspect_ratios_order:
            sizes += squares
            sizes += [(mn * math.sqrt(ar) / 2, mn / math.sqrt(ar) / 2) for ar in ratios if abs(ar - 1.0) >= 1e-6]
        else:
            sizes += [(mn * math.sqrt(ar) / 2, mn / math.sqrt(ar) / 2) for ar in ratios]
            sizes += squares[1:]
    dtype, device = input.dtype, input.device
    cx = (torch.arange(fw, dtype=dtype, device=device) + offset) * step_w
    cy = (torch.arange(fh, dtype=dtype, device=device) + offset) * step_h
    half = torch.as_tensor(sizes, dtype=dtype, device=device)           # [P, 2]
    cxg = cx[None, :, None].expand(fh, fw, len(sizes))
    cyg = cy[:, None, None].expand(fh, fw, len(sizes))
    boxes = torch.stack([(cxg - half[:, 0]) / iw, (cyg - half[:, 1]) / ih, (cxg + half[:, 0]) / iw, (cyg + half[:, 1]) / ih], -1)
    if clip:
        boxes = boxes.clamp(0.0, 1.0)
    var = torch.as_tensor(list(variance), dtype=dtype, device=device).expand(fh, fw, len(sizes), 4).contiguous()
    return boxes, var


def yolo_box(x, img_size, anchors, class_num, conf_thresh, downsample_ratio, clip_bbox=True, name=None, scale_x_y=1.0,
             iou_aware=False, iou_aware_factor=0.5):
    """Decode a YOLOv3 head ``x`` [N, A * (5 + C), H, W] (with ``iou_aware``: A extra IoU channels in front) into
    ``boxes`` [N, A*H*W, 4] in image pixels (``img_size`` [N, 2] = (height, width)) and ``scores`` [N, A*H*W, C]
    = objectness * class probability; cells whose objectness is below ``conf_thresh`` give zeros."""
    n, _, h, w = x.shape
    an = len(anchors) // 2
    if iou_aware:
        iou_pred, x = torch.sigmoid(x[:, :an]), x[:, an:]
    x = x.reshape(n, an, 5 + class_num, h, w)
    bias = -0.5 * (scale_x_y - 1.0)
    conf = torch.sigmoid(x[:, :, 4])
    if iou_aware:
        conf = conf.pow(1.0 - iou_aware_factor) * iou_pred.pow(iou_aware_factor)
    img_h = img_size[:, 0].to(x.dtype).reshape(n, 1, 1, 1)
    img_w = img_size[:, 1].to(x.dtype).reshape(n, 1, 1, 1)
    gx = torch.arange(w, dtype=x.dtype, device=x.device).reshape(1, 1, 1, w)
    gy = torch.arange(h, dtype=x.dtype, device=x.device).reshape(1, 1, h, 1)
    aw = torch.as_tensor(anchors[0::2], dtype=x.dtype, device=x.device).reshape(1, an, 1, 1)
    ah = torch.as_tensor(anchors[1::2], dtype=x.dtype, device=x.device).reshape(1, an, 1, 1)
    cx = (gx + torch.sigmoid(x[:, :, 0]) * scale_x_y + bias) * img_w / w
    cy = (gy + torch.sigmoid(x[:, :, 1]) * scale_x_y + bias) * img_h / h
    bw = torch.exp(x[:, :, 2]) * aw * img_w / (downsample_ratio * w)
    bh = torch.exp(x[:, :, 3]) * ah * img_h / (downsample_ratio * h)
    x1, y1, x2, y2 = cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2
    if clip_bbox:
        x1, y1 = x1.clamp_min(0), y1.clamp_min(0)
        x2, y2 = torch.minimum(x2, img_w - 1), torch.minimum(y2, img_h - 1)
    keep = (conf >= conf_thresh).to(x.dtype)
    boxes = torch.stack([x1, y1, x2, y2], -1) * keep[..., None]
    scores = (conf[:, :, None] * torch.sigmoid(x[:, :, 5:])) * keep[:, :, None]
    return boxes.reshape(n, an * h * w, 4), scores.permute(0, 1, 3, 4, 2).reshape(n, an * h * w, class_num)


def distribute_fpn_proposals(fpn_rois, min_level, max_level, refer_level, refer_scale, pixel_offset=False, rois_num=None,
                             name=None):
    """Send every RoI to the pyramid level floor(log2(sqrt(area) / refer_scale + 1e-6) + refer_level), clamped to
    [min_level, max_level].  Returns (per-level RoIs, restore index [R, 1] mapping the concatenated levels back to the
    input order, per-level per-image counts when ``rois_num`` is given)."""
    offset = 1.0 if pixel_offset else 0.0
    scale = torch.sqrt(_box_area(fpn_rois, offset))
    level = torch.floor(torch.log2(scale / refer_scale + 1e-6) + refer_level).clamp(min_level, max_level).to(torch.int64)
    image = _split_by_counts(fpn_rois, rois_num, 1) if rois_num is not None else None
    multi_rois, order, nums = [], [], []
    for lv in range(min_level, max_level + 1):
        idx = torch.nonzero(level == lv).reshape(-1)
        multi_rois.append(fpn_rois[idx])
        order.append(idx)
        if rois_num is not None:
            nums.append(torch.bincount(image[idx], minlength=rois_num.numel()).to(torch.int32))
    order = torch.cat(order)
    restore = torch.empty_like(order)
    restore[order] = torch.arange(order.numel(), device=order.device)
    return multi_rois, restore.reshape(-1, 1), (nums if rois_num is not None else None)


def generate_proposals(scores, bbox_deltas, img_size, anchors, variances, pre_nms_top_n=6000, post_nms_top_n=1000,
                       nms_thresh=0.5, min_size=0.1, eta=1.0, pixel_offset=False, return_rois_num=False, name=None):
    """RPN proposals.  Per image: take the ``pre_nms_top_n`` best anchors, decode their deltas (scaled by
    ``variances``; log-sizes clipped at log(1000 / 16)), clip to the image, drop boxes smaller than ``min_size``, run
    NMS and keep ``post_nms_top_n``.  ``scores`` [N, A, H, W], ``bbox_deltas`` [N, 4A, H, W], ``anchors`` /
    ``variances`` [H, W, A, 4], ``img_size`` [N, 2] = (height, width)."""
    n, a, h, w = scores.shape
    offset = 1.0 if pixel_offset else 0.0
    anchors = anchors.reshape(-1, 4)
    variances = variances.reshape(-1, 4)
    clip = math.log(1000.0 / 16.0)
    rois, probs, nums = [], [], []
    for b in range(n):
        s = scores[b].permute(1, 2, 0).reshape(-1)
        d = bbox_deltas[b].reshape(a, 4, h, w).permute(2, 3, 0, 1).reshape(-1, 4)
        order = torch.argsort(s, descending=True, stable=True)
        if 0 < pre_nms_top_n < order.numel():
            order = order[:pre_nms_top_n]
        s, d, an, var = s[order], d[order], anchors[order], variances[order]
        aw = an[:, 2] - an[:, 0] + offset
        ah = an[:, 3] - an[:, 1] + offset
        acx, acy = an[:, 0] + 0.5 * aw, an[:, 1] + 0.5 * ah
        cx = var[:, 0] * d[:, 0] * aw + acx
        cy = var[:, 1] * d[:, 1] * ah + acy
        bw = torch.exp(torch.clamp(var[:, 2] * d[:, 2], max=clip)) * aw
        bh = torch.exp(torch.clamp(var[:, 3] * d[:, 3], max=clip)) * ah
        box = torch.stack([cx - bw / 2, cy - bh / 2, cx + bw / 2 - offset, cy + bh / 2 - offset], 1)
        ih, iw = img_size[b, 0].to(box.dtype), img_size[b, 1].to(box.dtype)
        box = torch.stack([box[:, 0].clamp(min=0).minimum(iw - offset), box[:, 1].clamp(min=0).minimum(ih - offset),
                           box[:, 2].clamp(min=0).minimum(iw - offset), box[:, 3].clamp(min=0).minimum(ih - offset)], 1)
        ms = max(float(min_size), 1.0)
        ws, hs = box[:, 2] - box[:, 0] + offset, box[:, 3] - box[:, 1] + offset
        keep = (ws >= ms) & (hs >= ms)
        if pixel_offset:
            keep = keep & (box[:, 0] + ws / 2 <= iw) & (box[:, 1] + hs / 2 <= ih)
        box, s = box[keep], s[keep]
        kept = _greedy_nms(box, nms_thresh, offset, eta)
        if post_nms_top_n > 0:
            kept = kept[:post_nms_top_n]
        rois.append(box[kept])
        probs.append(s[kept].reshape(-1, 1))
        nums.append(kept.numel())
    rpn_rois, rpn_probs = torch.cat(rois), torch.cat(probs)
    if return_rois_num:
        return rpn_rois, rpn_probs, torch.as_tensor(nums, dtype=torch.int32, device=scores.device)
    return rpn_rois, rpn_probs, None


# ------------------------------------------------------------------------------------------------------------ RoI ops
def roi_pool(x, boxes, boxes_num, output_size, spatial_scale=1.0, name=None):
    """Max pooling of every RoI into ``output_size`` bins.  RoI corners are rounded to feature pixels
    (round(coord * spatial_scale)), the RoI is at least 1 x 1, bin (i, j) spans
    [floor(i * h / ph), ceil((i + 1) * h / ph)) of it, clipped to the map; an empty bin gives 0."""
    ph, pw = _pair(output_size)
    n, c, hh, ww = x.shape
    image = _split_by_counts(boxes, boxes_num, n)
    out = x.new_zeros((boxes.shape[0], c, ph, pw))
    scaled = [[_cround(v * spatial_scale) for v in row] for row in boxes.detach().cpu().tolist()]
    img_list = image.cpu().tolist()
    for r, ((x1, y1, x2, y2), b) in enumerate(zip(scaled, img_list)):
        rh, rw = max(y2 - y1 + 1, 1), max(x2 - x1 + 1, 1)
        for i in range(ph):
            hs = min(max(int(math.floor(i * rh / ph)) + y1, 0), hh)
            he = min(max(int(math.ceil((i + 1) * rh / ph)) + y1, 0), hh)
            for j in range(pw):
                ws = min(max(int(math.floor(j * rw / pw)) + x1, 0), ww)
                we = min(max(int(math.ceil((j + 1) * rw / pw)) + x1, 0), ww)
                if he > hs and we > ws:
                    out[r, :, i, j] = x[b, :, hs:he, ws:we].amax((-2, -1))
    return out


def psroi_pool(x, boxes, boxes_num, output_size, spatial_scale=1.0, name=None):
    """Position-sensitive RoI average pooling (R-FCN): output channel k at bin (i, j) averages INPUT channel
    (k * ph + i) * pw + j over that bin, so ``C`` must be ``out_channels * ph * pw``."""
    ph, pw = _pair(output_size)
    n, c, hh, ww = x.shape
    if c % (ph * pw) != 0:
        raise ValueError("the channels of input x should be divisible by output_size[0] * output_size[1]")
    oc = c // (ph * pw)
    image = _split_by_counts(boxes, boxes_num, n)
    out = x.new_zeros((boxes.shape[0], oc, ph, pw))
    coords = boxes.detach().cpu().tolist()
    img_list = image.cpu().tolist()
    xv = x.reshape(n, oc, ph, pw, hh, ww)
    for r, ((x1, y1, x2, y2), b) in enumerate(zip(coords, img_list)):
        sw, sh = _cround(x1) * spatial_scale, _cround(y1) * spatial_scale
        ew, eh = (_cround(x2) + 1.0) * spatial_scale, (_cround(y2) + 1.0) * spatial_scale
        rh, rw = max(eh - sh, 0.1), max(ew - sw, 0.1)
        bh, bw = rh / ph, rw / pw
        for i in range(ph):
            hs = min(max(int(math.floor(i * bh + sh)), 0), hh)
            he = min(max(int(math.ceil((i + 1) * bh + sh)), 0), hh)
            for j in range(pw):
                ws = min(max(int(math.floor(j * bw + sw)), 0), ww)
                we = min(max(int(math.ceil((j + 1) * bw + sw)), 0), ww)
                if he > hs and we > ws:
                    out[r, :, i, j] = xv[b, :, i, j, hs:he, ws:we].mean((-2, -1))
    return out


def _bilinear_gather(feat, y, x):
    """feat [C, H, W]; sample at (y, x) point lists with RoIAlign's border rule: points more than one pixel outside
    give 0, the rest are clamped into the map before interpolating."""
    c, h, w = feat.shape
    valid = (y >= -1.0) & (y <= h) & (x >= -1.0) & (x <= w)
    y = y.clamp(min=0)
    x = x.clamp(min=0)
    y0 = torch.floor(y).clamp(max=h - 1)
    x0 = torch.floor(x).clamp(max=w - 1)
    y1 = (y0 + 1).clamp(max=h - 1)
    x1 = (x0 + 1).clamp(max=w - 1)
    y = torch.where(y0 >= h - 1, y0, y)
    x = torch.where(x0 >= w - 1, x0, x)
    ly, lx = y - y0, x - x0
    hy, hx = 1 - ly, 1 - lx
    y0, y1, x0, x1 = y0.long(), y1.long(), x0.long(), x1.long()
    val = (feat[:, y0, x0] * (hy * hx) + feat[:, y0, x1] * (hy * lx) + feat[:, y1, x0] * (ly * hx) + feat[:, y1, x1] * (ly * lx))
    return val * valid.to(feat.dtype)


def roi_align(x, boxes, boxes_num, output_size, spatial_scale=1.0, sampling_ratio=-1, aligned=True, name=None):
    """Average of bilinear samples inside every bin (Mask R-CNN).  ``sampling_ratio`` points per bin side, or
    ceil(bin size) when <= 0; ``aligned`` shifts the box by half a pixel so that pixel centres sit at integers + 0.5
    (and lets RoIs be smaller than one pixel)."""
    ph, pw = _pair(output_size)
    n, c, hh, ww = x.shape
    image = _split_by_counts(boxes, boxes_num, n).cpu().tolist()
    out = []
    shift = 0.5 if aligned else 0.0
    for r in range(boxes.shape[0]):
        box = boxes[r] * spatial_scale - shift
        x1, y1, x2, y2 = box[0], box[1], box[2], box[3]
        rw, rh = x2 - x1, y2 - y1
        if not aligned:
            rw, rh = rw.clamp(min=1.0), rh.clamp(min=1.0)
        bh, bw = rh / ph, rw / pw
        gh = sampling_ratio if sampling_ratio > 0 else max(int(math.ceil(float(rh) / ph)), 1)
        gw = sampling_ratio if sampling_ratio > 0 else max(int(math.ceil(float(rw) / pw)), 1)
        iy = (torch.arange(ph * gh, dtype=x.dtype, device=x.device) + 0.5) / gh            # in units of bins
        ix = (torch.arange(pw * gw, dtype=x.dtype, device=x.device) + 0.5) / gw
        ys = (y1 + iy * bh)[:, None].expand(ph * gh, pw * gw).reshape(-1)
        xs = (x1 + ix * bw)[None, :].expand(ph * gh, pw * gw).reshape(-1)
        vals = _bilinear_gather(x[image[r]], ys, xs).reshape(c, ph, gh, pw, gw)
        out.append(vals.mean((2, 4)))
    return torch.stack(out) if out else x.new_zeros((0, c, ph, pw))


class RoIPool(Layer):
    def __init__(self, output_size, spatial_scale=1.0):
        super().__init__()
        self._output_size, self._spatial_scale = output_size, spatial_scale

    def forward(self, x, boxes, boxes_num):
        return roi_pool(x, boxes, boxes_num, self._output_size, self._spatial_scale)

    def extra_repr(self):
        return f"output_size={self._output_size}, spatial_scale={self._spatial_scale}"


class PSRoIPool(Layer):
    def __init__(self, output_size, spatial_scale=1.0):
        super().__init__()
        self.output_size, self.spatial_scale = output_size, spatial_scale

    def forward(self, x, boxes, boxes_num):
        return psroi_pool(x, boxes, boxes_num, self.output_size, self.spatial_scale)


class RoIAlign(Layer):
    def __init__(self, output_size, spatial_scale=1.0):
        super().__init__()
        self._output_size, self._spatial_scale = output_size, spatial_scale

    def forward(self, x, boxes, boxes_num, aligned=True):
        return roi_align(x, boxes, boxes_num, self._output_size, self._spatial_scale, aligned=aligned)


# ---------------------------------------------------------------------------------------------------- deformable conv
def deform_conv2d(x, offset, weight, bias=None, stride=1, padding=0, dilation=1, deformable_groups=1, groups=1, mask=None,
                  name=None):
    """Deformable convolution v1 (``mask=None``) / v2 (modulated).  Tap (i, j) of the kernel at output position p reads
    the input at p * stride - padding + (i, j) * dilation + offset, bilinearly, zero outside the map.

    ``offset`` [N, 2 * dg * kh * kw, Ho, Wo]: per deformable group and tap, (dy, dx).  ``mask`` [N, dg * kh * kw, Ho, Wo].
    Computed as an explicit im2col with bilinear gathers followed by a grouped matmul."""
    sh, sw = _pair(stride)
    ph, pw = _pair(padding)
    dh, dw = _pair(dilation)
    n, c, h, w = x.shape
    co, cg, kh, kw = weight.shape
    if c != cg * groups:
        raise ValueError(f"input channels {c} do not match weight {tuple(weight.shape)} with groups={groups}")
    ho = (h + 2 * ph - (dh * (kh - 1) + 1)) // sh + 1
    wo = (w + 2 * pw - (dw * (kw - 1) + 1)) // sw + 1
    dg, k = deformable_groups, kh * kw
    if tuple(offset.shape) != (n, 2 * dg * k, ho, wo):
        raise ValueError(f"offset should be {(n, 2 * dg * k, ho, wo)}, got {tuple(offset.shape)}")
    if c % dg != 0:
        raise ValueError("input channels must be divisible by deformable_groups")
    off = offset.reshape(n, dg, k, 2, ho, wo)
    base_y = (torch.arange(ho, device=x.device, dtype=x.dtype) * sh - ph).reshape(1, 1, 1, ho, 1)
    base_x = (torch.arange(wo, device=x.device, dtype=x.dtype) * sw - pw).reshape(1, 1, 1, 1, wo)
    tap_y = (torch.arange(kh, device=x.device, dtype=x.dtype) * dh).repeat_interleave(kw).reshape(1, 1, k, 1, 1)
    tap_x = (torch.arange(kw, device=x.device, dtype=x.dtype) * dw).repeat(kh).reshape(1, 1, k, 1, 1)
    py = base_y + tap_y + off[:, :, :, 0]                                  # [n, dg, k, ho, wo]
    px = base_x + tap_x + off[:, :, :, 1]
    y0, x0 = torch.floor(py), torch.floor(px)
    ly, lx = py - y0, px - x0
    cpg = c // dg
    xg = x.reshape(n, dg, cpg, h * w)

    def corner(yy, xx, wgt):
        inside = ((yy >= 0) & (yy <= h - 1) & (xx >= 0) & (xx <= w - 1)).to(x.dtype)
        idx = (yy.clamp(0, h - 1) * w + xx.clamp(0, w - 1)).long().reshape(n, dg, 1, -1).expand(n, dg, cpg, k * ho * wo)
        vals = torch.gather(xg, 3, idx).reshape(n, dg, cpg, k, ho, wo)
        return vals * (wgt * inside).unsqueeze(2)
    cols = (corner(y0, x0, (1 - ly) * (1 - lx)) + corner(y0, x0 + 1, (1 - ly) * lx)
            + corner(y0 + 1, x0, ly * (1 - lx)) + corner(y0 + 1, x0 + 1, ly * lx))
    if mask is not None:
        if tuple(mask.shape) != (n, dg * k, ho, wo):
            raise ValueError(f"mask should be {(n, dg * k, ho, wo)}, got {tuple(mask.shape)}")
        cols = cols * mask.reshape(n, dg, 1, k, ho, wo)
    cols = cols.reshape(n, groups, cg * k, ho * wo)                       # channels are (dg, cpg) = c in order
    wmat = weight.reshape(groups, co // groups, cg * k)
    out = torch.einsum("gok,ngkp->ngop", wmat, cols).reshape(n, co, ho, wo)
    if bias is not None:
        out = out + bias.reshape(1, -1, 1, 1)
    return out


class DeformConv2D(Layer):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, deformable_groups=1,
                 groups=1, weight_attr=None, bias_attr=None):
        super().__init__()
        assert weight_attr is not False, "weight_attr should not be False in Conv."
        if in_channels % groups != 0:
            raise ValueError("in_channels must be divisible by groups.")
        self._in_channels, self._out_channels, self._groups = in_channels, out_channels, groups
        self._deformable_groups = deformable_groups
        self._kernel_size, self._stride, self._padding, self._dilation = _pair(kernel_size), _pair(stride), _pair(padding), _pair(dilation)
        shape = [out_channels, in_channels // groups, *self._kernel_size]
        std = math.sqrt(2.0 / (math.prod(self._kernel_size) * in_channels))
        self.weight = self.create_parameter(shape, attr=weight_attr, default_initializer=I.Normal(0.0, std))
        self.bias = self.create_parameter([out_channels], attr=bias_attr, is_bias=True)

    def forward(self, x, offset, mask=None):
        return deform_conv2d(x, offset, self.weight, self.bias, self._stride, self._padding, self._dilation,
                             self._deformable_groups, self._groups, mask)


# ------------------------------------------------------------------------------------------------------------ file I/O
def read_file(filename, name=None):
    """The bytes of a file as a 1-D uint8 tensor."""
    with open(filename, "rb") as f:
        return torch.frombuffer(bytearray(f.read()), dtype=torch.uint8)


def decode_jpeg(x, mode="unchanged", name=None):
    """Encoded JPEG bytes (1-D uint8 tensor) -> CHW uint8 image.  ``mode``: "unchanged", "gray" or "rgb"."""
    import io

    from PIL import Image
    img = Image.open(io.BytesIO(bytes(x.cpu().numpy().tobytes())))
    mode = mode.lower()
    if mode == "gray":
        img = img.convert("L")
    elif mode == "rgb":
        img = img.convert("RGB")
    elif mode != "unchanged":
        raise ValueError("mode should be 'unchanged', 'gray' or 'rgb'")
    arr = np.asarray(img)
    if arr.ndim == 2:
        arr = arr[:, :, None]
    return torch.from_numpy(np.array(arr.transpose(2, 0, 1), copy=True, order="C"))


# ----------------------------------------------------------------------------------------------------------- YOLO loss
def _sce(logit, label):
    """Sigmoid cross entropy on logits, the overflow-free form."""
    return logit.clamp_min(0) - logit * label + torch.log1p(torch.exp(-logit.abs()))


def _cxcywh_iou(b1, b2):
    """IoU of centre-size boxes, broadcasting; no clamping of degenerate unions (inputs are valid boxes)."""
    def overlap(c1, w1, c2, w2):
        return torch.minimum(c1 + w1 / 2, c2 + w2 / 2) - torch.maximum(c1 - w1 / 2, c2 - w2 / 2)
    ow = overlap(b1[..., 0], b1[..., 2], b2[..., 0], b2[..., 2])
    oh = overlap(b1[..., 1], b1[..., 3], b2[..., 1], b2[..., 3])
    inter = torch.where((ow < 0) | (oh < 0), torch.zeros_like(ow), ow * oh)
    return inter / (b1[..., 2] * b1[..., 3] + b2[..., 2] * b2[..., 3] - inter)


def yolo_loss(x, gt_box, gt_label, anchors, anchor_mask, class_num, ignore_thresh, downsample_ratio, gt_score=None,
              use_label_smooth=True, name=None, scale_x_y=1.0):
    """YOLOv3 loss of one head, one value per image.

    ``x`` [N, M * (5 + C), H, W] for the M anchors of ``anchor_mask``; ``gt_box`` [N, B, 4] = normalised (cx, cy, w, h),
    boxes with w or h <= 0 are padding; ``gt_label`` [N, B].  Each ground truth is matched to the anchor (of ALL
    anchors) whose shape overlaps it best; when that anchor belongs to this head, the cell containing the box centre
    is responsible: sigmoid-CE on the centre offsets, L1 on log sizes (both weighted by 2 - w*h), sigmoid-CE class
    loss, objectness target = ``gt_score``.  Cells that predict a box overlapping some ground truth by more than
    ``ignore_thresh`` and are not responsible are left out of the objectness loss; all others are negatives."""
    n, _, h, w = x.shape
    mask_num, an_num = len(anchor_mask), len(anchors) // 2
    b = gt_box.shape[1]
    input_size = downsample_ratio * h
    bias = -0.5 * (scale_x_y - 1.0)
    xr = x.reshape(n, mask_num, 5 + class_num, h, w)
    pos, neg = 1.0, 0.0
    if use_label_smooth:
        smooth = min(1.0 / class_num, 1.0 / 40)
        pos, neg = 1.0 - smooth, smooth
    score = gt_score if gt_score is not None else torch.ones((n, b), dtype=x.dtype, device=x.device)
    with torch.no_grad():
        valid = (gt_box[..., 2] > 1e-6) & (gt_box[..., 3] > 1e-6)                                   # [n, b]
        # predicted boxes of this head, normalised as the reference does (grid size h for both axes)
        gx = torch.arange(w, dtype=x.dtype, device=x.device).reshape(1, 1, 1, w)
        gy = torch.arange(h, dtype=x.dtype, device=x.device).reshape(1, 1, h, 1)
        aw = torch.as_tensor([anchors[2 * m] for m in anchor_mask], dtype=x.dtype, device=x.device).reshape(1, mask_num, 1, 1)
        ah = torch.as_tensor([anchors[2 * m + 1] for m in anchor_mask], dtype=x.dtype, device=x.device).reshape(1, mask_num, 1, 1)
        pred = torch.stack([(gx + torch.sigmoid(xr[:, :, 0]) * scale_x_y + bias) / h, (gy + torch.sigmoid(xr[:, :, 1]) * scale_x_y + bias) / h,
                            torch.exp(xr[:, :, 2]) * aw / input_size, torch.exp(xr[:, :, 3]) * ah / input_size], -1)
        iou = _cxcywh_iou(pred[:, :, :, :, None, :], gt_box[:, None, None, None, :, :])               # [n, m, h, w, b]
        iou = torch.where(valid[:, None, None, None, :], iou, torch.zeros_like(iou))
        best = iou.max(-1).values if b > 0 else torch.zeros_like(pred[..., 0])
        obj_mask = torch.where(best > ignore_thresh, -torch.ones_like(best), torch.zeros_like(best))   # [n, m, h, w]
        # anchor matching by shape
        all_aw = torch.as_tensor(anchors[0::2], dtype=x.dtype, device=x.device) / input_size
        all_ah = torch.as_tensor(anchors[1::2], dtype=x.dtype, device=x.device) / input_size
        zeros = torch.zeros_like(all_aw)
        an_boxes = torch.stack([zeros, zeros, all_aw, all_ah], -1)                                     # [A, 4]
        shifted = torch.cat([torch.zeros_like(gt_box[..., :2]), gt_box[..., 2:]], -1)
        an_iou = _cxcywh_iou(an_boxes[None, None], shifted[:, :, None, :])                             # [n, b, A]
        best_n = an_iou.argmax(-1)                                                                     # first maximum
        lut = torch.full((an_num,), -1, dtype=torch.int64, device=x.device)
        for i, m in enumerate(anchor_mask):
            lut[m] = i
        match = torch.where(valid, lut[best_n], torch.full_like(best_n, -1))                           # [n, b]
        gi = (gt_box[..., 0] * w).to(torch.int64).clamp(0, w - 1)
        gj = (gt_box[..., 1] * h).to(torch.int64).clamp(0, h - 1)
    loss = x.new_zeros(n)
    sel = torch.nonzero(match >= 0)                                                                    # rows (image, gt)
    if sel.numel():
        bi, ti = sel[:, 0], sel[:, 1]
        mi, cj, ci = match[bi, ti], gj[bi, ti], gi[bi, ti]
        gt = gt_box[bi, ti]
        sc = score[bi, ti]
        cell = xr[bi, mi, :, cj, ci]                                                                   # [k, 5 + C]
        an_w = torch.as_tensor(anchors[0::2], dtype=x.dtype, device=x.device)[best_n[bi, ti]]
        an_h = torch.as_tensor(anchors[1::2], dtype=x.dtype, device=x.device)[best_n[bi, ti]]
        tx, ty = gt[:, 0] * h - ci.to(x.dtype), gt[:, 1] * h - cj.to(x.dtype)
        tw, th = torch.log(gt[:, 2] * input_size / an_w), torch.log(gt[:, 3] * input_size / an_h)
        scale = (2.0 - gt[:, 2] * gt[:, 3]) * sc
        loc = (_sce(cell[:, 0], tx) + _sce(cell[:, 1], ty) + (cell[:, 2] - tw).abs() + (cell[:, 3] - th).abs()) * scale
        target = torch.full((sel.shape[0], class_num), neg, dtype=x.dtype, device=x.device)
        target[torch.arange(sel.shape[0], device=x.device), gt_label[bi, ti].long()] = pos
        cls = _sce(cell[:, 5:], target).sum(-1) * sc
        loss = loss.index_add(0, bi, loc + cls)
        with torch.no_grad():
            # later ground truths overwrite earlier ones in the same cell, as the sequential reference does
            for r in range(sel.shape[0]):
                obj_mask[bi[r], mi[r], cj[r], ci[r]] = sc[r]
    obj_logit = xr[:, :, 4]
    positive = (obj_mask > 1e-5).to(x.dtype)
    negative = ((obj_mask <= 1e-5) & (obj_mask > -0.5)).to(x.dtype)
    obj = _sce(obj_logit, torch.ones_like(obj_logit)) * obj_mask.clamp_min(0) * positive + _sce(obj_logit, torch.zeros_like(obj_logit)) * negative
    return loss + obj.sum((1, 2, 3))


_DEFAULT = object()


class ConvNormActivation(torch.nn.Sequential):
    """Conv2D + norm layer + activation, the block the classification models are built from."""

    def __init__(self, in_channels, out_channels, kernel_size=3, stride=1, padding=None, groups=1, norm_layer=_DEFAULT,
                 activation_layer=_DEFAULT, dilation=1, bias=None):
        from .. import nn
        norm_layer = nn.BatchNorm2D if norm_layer is _DEFAULT else norm_layer          # None: no norm layer
        activation_layer = nn.ReLU if activation_layer is _DEFAULT else activation_layer
        if padding is None:
            padding = (kernel_size - 1) // 2 * dilation
        if bias is None:
            bias = norm_layer is None
        layers = [nn.Conv2D(in_channels, out_channels, kernel_size, stride, padding, dilation=dilation, groups=groups,
                            bias_attr=None if bias else False)]
        if norm_layer is not None:
            layers.append(norm_layer(out_channels))
        if activation_layer is not None:
            layers.append(activation_layer())
        super().__init__(*layers)
