"""paddle.vision.ops against torchvision's operators and loop restatements of the reference kernels
(python/paddle/vision/ops.py, paddle/phi/kernels/cpu/{yolo_loss,yolo_box,prior_box,matrix_nms,...}_kernel.cc)."""
import io
import math

import numpy as np
import pytest
import torch

import paddle
import paddle.vision.ops as O

tv = pytest.importorskip("torchvision.ops")


def _boxes(n, size=50.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    xy = torch.rand(n, 2, generator=g) * size
    wh = torch.rand(n, 2, generator=g) * size / 2 + 1
    return torch.cat([xy, xy + wh], 1)


def test_nms_variants():
    boxes, scores = _boxes(60), torch.rand(60, generator=torch.Generator().manual_seed(1))
    for thr in (0.1, 0.3, 0.6):
        torch.testing.assert_close(O.nms(boxes, thr, scores), tv.nms(boxes, scores, thr))
        order = scores.argsort(descending=True)
        torch.testing.assert_close(O.nms(boxes[order], thr), tv.nms(boxes[order], scores[order], thr))   # given order
    cats = torch.randint(0, 3, (60,), generator=torch.Generator().manual_seed(2))
    got = O.nms(boxes, 0.3, scores, cats, [0, 1, 2])
    want = tv.batched_nms(boxes, scores, cats, 0.3)
    torch.testing.assert_close(got, want)
    assert O.nms(boxes, 0.3, scores, cats, [0, 1, 2], top_k=5).tolist() == want[:5].tolist()
    assert O.nms(boxes, 0.3, scores, cats, [1]).tolist() == [i for i in want.tolist() if cats[i] == 1]
    assert O.nms(torch.zeros(0, 4), 0.5).numel() == 0


def test_matrix_nms_matches_sequential_restatement():
    torch.manual_seed(0)
    boxes = _boxes(12, seed=3)[None] / 80.0
    scores = torch.rand(1, 3, 12)

    def restate(cls_scores, bx, gaussian):
        keep = [i for i in np.argsort(-cls_scores.numpy(), kind="stable") if cls_scores[i] > 0.1]
        iou = O._iou_matrix(bx[keep], bx[keep]).numpy()
        out = []
        for i, idx in enumerate(keep):
            decay = 1.0
            for j in range(i):
                comp = max([iou[j, k] for k in range(j)], default=0.0)
                d = math.exp((comp ** 2 - iou[i, j] ** 2) * 2.0) if gaussian else (1 - iou[i, j]) / (1 - comp)
                decay = min(decay, d)
            s = float(cls_scores[idx]) * decay
            if s > 0.05:
                out.append((s, idx))
        return out
    for gaussian in (False, True):
        out, index, num = O.matrix_nms(boxes, scores, 0.1, 0.05, -1, -1, gaussian, 2.0, background_label=0, return_index=True)
        want = sorted([(s, idx, c) for c in (1, 2) for s, idx in restate(scores[0, c], boxes[0], gaussian)], key=lambda t: -t[0])
        assert int(num[0]) == len(want) == out.shape[0]
        np.testing.assert_allclose(out[:, 1].numpy(), [w[0] for w in want], rtol=1e-5)
        assert index.reshape(-1).tolist() == [w[1] for w in want] and out[:, 0].tolist() == [float(w[2]) for w in want]
        torch.testing.assert_close(out[:, 2:], boxes[0][index.reshape(-1)])
    top = O.matrix_nms(boxes, scores, 0.1, 0.05, 4, 3, return_rois_num=False)
    assert top.shape == (3, 6)
    empty, num = O.matrix_nms(boxes, scores, 2.0, 0.05, -1, -1)
    assert empty.shape == (0, 6) and num.tolist() == [0]


def test_roi_ops_match_torchvision():
    torch.manual_seed(0)
    x = torch.randn(2, 8, 16, 20, dtype=torch.float64, requires_grad=True)
    rois = torch.tensor([[1.3, 2.1, 9.7, 12.4], [0.0, 0.0, 19.0, 15.0], [4.2, 3.3, 6.1, 5.0], [10.5, 1.0, 18.2, 14.9], [-3.0, -2.0, 4.0, 5.0]],
                        dtype=torch.float64)
    num = torch.tensor([3, 2], dtype=torch.int32)
    tv_rois = torch.cat([torch.tensor([0., 0, 0, 1, 1], dtype=torch.float64)[:, None], rois], 1)
    for aligned in (True, False):
        for ratio in (2, -1):
            for scale in (1.0, 0.5):
                got = O.roi_align(x, rois, num, (3, 4), scale, ratio, aligned)
                want = tv.roi_align(x, tv_rois, (3, 4), scale, ratio, aligned)
                torch.testing.assert_close(got, want, rtol=1e-9, atol=1e-9)
    g1, = torch.autograd.grad(O.roi_align(x, rois, num, 3, 0.5, 2).square().sum(), x)
    g2, = torch.autograd.grad(tv.roi_align(x, tv_rois, 3, 0.5, 2, True).square().sum(), x)
    torch.testing.assert_close(g1, g2, rtol=1e-8, atol=1e-8)
    assert O.RoIAlign(2, 0.25)(x, rois, num).shape == (5, 8, 2, 2)
    int_rois = torch.tensor([[1., 2, 9, 12], [0, 0, 19, 15], [4, 3, 6, 5], [10, 1, 18, 14], [3, 3, 3, 3]], dtype=torch.float64)
    tv_int = torch.cat([tv_rois[:, :1], int_rois], 1)
    for scale in (1.0, 0.5):
        torch.testing.assert_close(O.roi_pool(x, int_rois, num, (3, 2), scale), tv.roi_pool(x, tv_int, (3, 2), scale))
    g1, = torch.autograd.grad(O.roi_pool(x, int_rois, num, 2).sum(), x)
    g2, = torch.autograd.grad(tv.roi_pool(x, tv_int, 2).sum(), x)
    torch.testing.assert_close(g1, g2)
    xp = torch.randn(2, 2 * 3 * 3, 16, 20, dtype=torch.float64)
    img = [0, 0, 0, 1, 1]
    for scale in (1.0, 0.5):
        # the reference's rule: RoI = [round(x1), round(x2) + 1) * scale, bins floor / ceil, clipped, empty bins 0
        want = np.zeros((5, 2, 3, 3))
        xn = xp.numpy()
        for r, (x1, y1, x2, y2) in enumerate(int_rois.tolist()):
            sw, sh, ew, eh = x1 * scale, y1 * scale, (x2 + 1) * scale, (y2 + 1) * scale
            bh, bw = max(eh - sh, 0.1) / 3, max(ew - sw, 0.1) / 3
            for c in range(2):
                for i in range(3):
                    for j in range(3):
                        hs, he = min(max(math.floor(i * bh + sh), 0), 16), min(max(math.ceil((i + 1) * bh + sh), 0), 16)
                        ws, we = min(max(math.floor(j * bw + sw), 0), 20), min(max(math.ceil((j + 1) * bw + sw), 0), 20)
                        if he > hs and we > ws:
                            want[r, c, i, j] = xn[img[r], (c * 3 + i) * 3 + j, hs:he, ws:we].mean()
        np.testing.assert_allclose(O.psroi_pool(xp, int_rois, num, 3, scale).numpy(), want, rtol=1e-9, atol=1e-12)
    assert O.PSRoIPool(3)(xp, int_rois, num).shape == (5, 2, 3, 3) and O.RoIPool(2)(x, int_rois, num).shape == (5, 8, 2, 2)
    with pytest.raises(ValueError):
        O.psroi_pool(x, int_rois, num, 3)
    with pytest.raises(ValueError):
        O.roi_pool(x, int_rois, torch.tensor([1, 1]), 2)


def test_deform_conv2d_matches_torchvision():
    torch.manual_seed(0)
    for groups, dg, use_mask in ((1, 1, False), (2, 1, True), (1, 2, True), (2, 4, False)):
        x = torch.randn(2, 8, 9, 10, dtype=torch.float64, requires_grad=True)
        w = torch.randn(6, 8 // groups, 3, 2, dtype=torch.float64, requires_grad=True)
        bias = torch.randn(6, dtype=torch.float64)
        stride, pad, dil = (2, 1), (1, 2), (1, 2)
        ho = (9 + 2 - (1 * 2 + 1)) // 2 + 1
        wo = (10 + 4 - (2 * 1 + 1)) // 1 + 1
        off = torch.randn(2, 2 * dg * 6, ho, wo, dtype=torch.float64, requires_grad=True) * 1.5
        mask = torch.rand(2, dg * 6, ho, wo, dtype=torch.float64) if use_mask else None
        got = O.deform_conv2d(x, off, w, bias, stride, pad, dil, dg, groups, mask)
        want = tv.deform_conv2d(x, off, w, bias, stride, pad, dil, mask)
        torch.testing.assert_close(got, want, rtol=1e-9, atol=1e-9)
        ga = torch.autograd.grad(got.square().sum(), [x, w, off])
        gb = torch.autograd.grad(want.square().sum(), [x, w, off])
        for a, b in zip(ga, gb):
            torch.testing.assert_close(a, b, rtol=1e-7, atol=1e-8)
    # zero offsets: an ordinary convolution
    x, w = torch.randn(1, 4, 7, 7), torch.randn(5, 4, 3, 3)
    torch.testing.assert_close(O.deform_conv2d(x, torch.zeros(1, 18, 7, 7), w, padding=1), torch.nn.functional.conv2d(x, w, padding=1),
                               rtol=1e-4, atol=1e-5)
    layer = O.DeformConv2D(4, 5, 3, padding=1, deformable_groups=2)
    assert layer.weight.shape == (5, 4, 3, 3) and layer(x, torch.zeros(1, 36, 7, 7), torch.ones(1, 18, 7, 7)).shape == (1, 5, 7, 7)
    with pytest.raises(ValueError):
        O.deform_conv2d(x, torch.zeros(1, 10, 7, 7), w, padding=1)


def test_box_coder_prior_box_fpn_and_proposals():
    torch.manual_seed(0)
    priors, targets = _boxes(5, seed=4), _boxes(3, seed=5)
    var = [0.1, 0.1, 0.2, 0.2]
    for normalized in (True, False):
        norm = 0.0 if normalized else 1.0
        enc = O.box_coder(priors, var, targets, "encode_center_size", normalized)
        assert enc.shape == (3, 5, 4)
        pw, ph = priors[:, 2] - priors[:, 0] + norm, priors[:, 3] - priors[:, 1] + norm
        pcx, pcy = priors[:, 0] + pw / 2, priors[:, 1] + ph / 2
        t = targets[1]
        tw, th = t[2] - t[0] + norm, t[3] - t[1] + norm
        want = torch.stack([((t[0] + t[2]) / 2 - pcx[2]) / pw[2] / 0.1, ((t[1] + t[3]) / 2 - pcy[2]) / ph[2] / 0.1,
                            torch.log(tw / pw[2]) / 0.2, torch.log(th / ph[2]) / 0.2])
        torch.testing.assert_close(enc[1, 2], want)
        # decoding the code of a box against the same priors gives the box back (up to the centre convention)
        dec = O.box_coder(priors, var, enc, "decode_center_size", normalized, axis=0)
        cx = (targets[:, 0] + targets[:, 2]) / 2
        torch.testing.assert_close((dec[:, 0, 0] + dec[:, 0, 2] + norm) / 2, cx, rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(dec[:, :, 2] - dec[:, :, 0] + norm, (targets[:, 2] - targets[:, 0] + norm)[:, None].expand(3, 5), rtol=1e-4, atol=1e-4)
    var_t = torch.tensor(var).expand(5, 4).contiguous()
    torch.testing.assert_close(O.box_coder(priors, var_t, targets), O.box_coder(priors, var, targets))
    torch.testing.assert_close(O.box_coder(priors, None, targets) / torch.tensor(var), O.box_coder(priors, var, targets))
    codes = torch.randn(5, 3, 4) * 0.1
    a1 = O.box_coder(priors, var, codes, "decode_center_size", axis=1)                 # priors along dim 0
    a0 = O.box_coder(priors, var, codes.transpose(0, 1).contiguous(), "decode_center_size", axis=0)
    torch.testing.assert_close(a1, a0.transpose(0, 1))
    with pytest.raises(ValueError):
        O.box_coder(priors, var, targets, "encode")
    # prior boxes: count, order and geometry
    feat, img = torch.zeros(1, 8, 3, 4), torch.zeros(1, 3, 30, 40)
    boxes, variances = O.prior_box(feat, img, min_sizes=[4.0, 8.0], max_sizes=[8.0, 16.0], aspect_ratios=[2.0], flip=True, clip=True)
    assert boxes.shape == (3, 4, 8, 4) and variances.shape == (3, 4, 8, 4)             # per min size: ratios 1, 2, 1/2 + one max square
    torch.testing.assert_close(variances[1, 2, 5], torch.tensor([0.1, 0.1, 0.2, 0.2]))
    cx, cy = (2 + 0.5) * 10.0, (1 + 0.5) * 10.0
    def box(hw, hh):
        return torch.tensor([(cx - hw) / 40, (cy - hh) / 30, (cx + hw) / 40, (cy + hh) / 30])
    torch.testing.assert_close(boxes[1, 2, 0], box(2.0, 2.0))
    torch.testing.assert_close(boxes[1, 2, 1], box(4 * math.sqrt(2) / 2, 4 / math.sqrt(2) / 2))
    torch.testing.assert_close(boxes[1, 2, 2], box(4 / math.sqrt(2) / 2, 4 * math.sqrt(2) / 2))
    torch.testing.assert_close(boxes[1, 2, 3], box(math.sqrt(32.0) / 2, math.sqrt(32.0) / 2))
    torch.testing.assert_close(boxes[1, 2, 4], box(4.0, 4.0))
    caffe, _ = O.prior_box(feat, img, min_sizes=[4.0], max_sizes=[8.0], aspect_ratios=[2.0], min_max_aspect_ratios_order=True)
    torch.testing.assert_close(caffe[1, 2, 1], box(math.sqrt(32.0) / 2, math.sqrt(32.0) / 2))      # max square second
    stepped, _ = O.prior_box(feat, img, min_sizes=[4.0], steps=[8.0, 8.0], offset=0.0)
    torch.testing.assert_close(stepped[2, 3, 0], torch.tensor([(24 - 2) / 40, (16 - 2) / 30, (24 + 2) / 40, (16 + 2) / 30]))
    assert float(boxes.min()) >= 0 and float(boxes.max()) <= 1
    # FPN distribution
    rois = torch.tensor([[0., 0, 10, 10], [0, 0, 100, 100], [0, 0, 300, 300], [5, 5, 60, 60], [0, 0, 1000, 1000]])
    multi, restore, nums = O.distribute_fpn_proposals(rois, 2, 5, 4, 224, rois_num=torch.tensor([3, 2]))
    levels = [int(min(max(math.floor(math.log2(math.sqrt(float((r[2] - r[0]) * (r[3] - r[1]))) / 224 + 1e-6) + 4), 2), 5)) for r in rois]
    assert [m.shape[0] for m in multi] == [levels.count(lv) for lv in range(2, 6)]
    torch.testing.assert_close(torch.cat(multi)[restore.reshape(-1)], rois)
    assert [n.tolist() for n in nums] == [[sum(1 for i in range(3) if levels[i] == lv), sum(1 for i in range(3, 5) if levels[i] == lv)] for lv in range(2, 6)]
    assert O.distribute_fpn_proposals(rois, 2, 5, 4, 224)[2] is None
    # RPN proposals against a per-image restatement through torchvision's nms
    n, a, h, w = 2, 3, 4, 5
    scores = torch.rand(n, a, h, w)
    deltas = torch.randn(n, 4 * a, h, w) * 0.3
    anchors = _boxes(h * w * a, size=40.0, seed=6).reshape(h, w, a, 4)
    variances = torch.full((h, w, a, 4), 0.5)
    img_size = torch.tensor([[48.0, 64.0], [40.0, 50.0]])
    rois, probs, num = O.generate_proposals(scores, deltas, img_size, anchors, variances, pre_nms_top_n=40, post_nms_top_n=10,
                                            nms_thresh=0.5, min_size=2.0, return_rois_num=True)
    assert rois.shape[0] == int(num.sum()) == probs.shape[0] and all(int(k) <= 10 for k in num)
    start = 0
    for b in range(n):
        s = scores[b].permute(1, 2, 0).reshape(-1)
        d = deltas[b].reshape(a, 4, h, w).permute(2, 3, 0, 1).reshape(-1, 4)
        order = s.argsort(descending=True)[:40]
        an = anchors.reshape(-1, 4)[order]
        aw, ah = an[:, 2] - an[:, 0], an[:, 3] - an[:, 1]
        cx, cy = 0.5 * d[order, 0] * aw + an[:, 0] + aw / 2, 0.5 * d[order, 1] * ah + an[:, 1] + ah / 2
        bw, bh = torch.exp(0.5 * d[order, 2]) * aw, torch.exp(0.5 * d[order, 3]) * ah
        bx = torch.stack([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2], 1)
        bx[:, 0::2] = bx[:, 0::2].clamp(0, float(img_size[b, 1]))
        bx[:, 1::2] = bx[:, 1::2].clamp(0, float(img_size[b, 0]))
        ok = ((bx[:, 2] - bx[:, 0]) >= 2.0) & ((bx[:, 3] - bx[:, 1]) >= 2.0)
        bx, sc = bx[ok], s[order][ok]
        keep = tv.nms(bx, sc, 0.5)[:10]
        k = int(num[b])
        assert k == keep.numel()
        torch.testing.assert_close(rois[start:start + k], bx[keep])
        torch.testing.assert_close(probs[start:start + k, 0], sc[keep])
        start += k
    assert O.generate_proposals(scores, deltas, img_size, anchors, variances)[2] is None


def _yolo_loss_loops(x, gt_box, gt_label, gt_score, anchors, anchor_mask, class_num, ignore_thresh, downsample, smooth, scale_xy):
    """The reference kernel's loops (yolo_loss_kernel.cc:180-370) in numpy."""
    n, _, h, w = x.shape
    m, b = len(anchor_mask), gt_box.shape[1]
    input_size = downsample * h
    bias = -0.5 * (scale_xy - 1.0)
    xr = x.reshape(n, m, 5 + class_num, h, w)
    sig = lambda v: 1 / (1 + np.exp(-v))                                                  # noqa: E731
    sce = lambda v, t: max(v, 0.0) - v * t + math.log(1 + math.exp(-abs(v)))              # noqa: E731
    pos, neg = (1.0, 0.0)
    if smooth:
        sw = min(1.0 / class_num, 1.0 / 40)
        pos, neg = 1 - sw, sw

    def iou(b1, b2):
        def ov(c1, w1, c2, w2):
            return min(c1 + w1 / 2, c2 + w2 / 2) - max(c1 - w1 / 2, c2 - w2 / 2)
        ow, oh = ov(b1[0], b1[2], b2[0], b2[2]), ov(b1[1], b1[3], b2[1], b2[3])
        inter = 0.0 if ow < 0 or oh < 0 else ow * oh
        return inter / (b1[2] * b1[3] + b2[2] * b2[3] - inter)
    loss = np.zeros(n)
    obj = np.zeros((n, m, h, w))
    for i in range(n):
        valid = [gt_box[i, t, 2] > 1e-6 and gt_box[i, t, 3] > 1e-6 for t in range(b)]
        for j in range(m):
            for k in range(h):
                for l in range(w):
                    an = anchor_mask[j]
                    pred = [(l + sig(xr[i, j, 0, k, l]) * scale_xy + bias) / h, (k + sig(xr[i, j, 1, k, l]) * scale_xy + bias) / h,
                            math.exp(xr[i, j, 2, k, l]) * anchors[2 * an] / input_size, math.exp(xr[i, j, 3, k, l]) * anchors[2 * an + 1] / input_size]
                    best = max([iou(pred, gt_box[i, t]) for t in range(b) if valid[t]], default=0.0)
                    if best > ignore_thresh:
                        obj[i, j, k, l] = -1
        for t in range(b):
            if not valid[t]:
                continue
            gt = gt_box[i, t]
            gi, gj = int(gt[0] * w), int(gt[1] * h)
            best_iou, best_n = 0.0, 0
            for an in range(len(anchors) // 2):
                v = iou([0, 0, anchors[2 * an] / input_size, anchors[2 * an + 1] / input_size], [0, 0, gt[2], gt[3]])
                if v > best_iou:
                    best_iou, best_n = v, an
            if best_n not in anchor_mask:
                continue
            mi = anchor_mask.index(best_n)
            sc = gt_score[i, t]
            cell = xr[i, mi, :, gj, gi]
            scale = (2.0 - gt[2] * gt[3]) * sc
            loss[i] += (sce(cell[0], gt[0] * h - gi) + sce(cell[1], gt[1] * h - gj)
                        + abs(cell[2] - math.log(gt[2] * input_size / anchors[2 * best_n]))
                        + abs(cell[3] - math.log(gt[3] * input_size / anchors[2 * best_n + 1]))) * scale
            obj[i, mi, gj, gi] = sc
            for c in range(class_num):
                loss[i] += sce(cell[5 + c], pos if c == gt_label[i, t] else neg) * sc
        for j in range(m):
            for k in range(h):
                for l in range(w):
                    o = obj[i, j, k, l]
                    if o > 1e-5:
                        loss[i] += sce(xr[i, j, 4, k, l], 1.0) * o
                    elif o > -0.5:
                        loss[i] += sce(xr[i, j, 4, k, l], 0.0)
    return loss


def test_yolo_box_and_loss():
    torch.manual_seed(0)
    anchors, class_num = [10, 13, 16, 30, 33, 23], 4
    x = torch.randn(2, 3 * (5 + class_num), 5, 5, dtype=torch.float64)
    img_size = torch.tensor([[160, 200], [120, 96]], dtype=torch.int32)
    boxes, scores = O.yolo_box(x, img_size, anchors, class_num, 0.3, 32, clip_bbox=True, scale_x_y=1.2)
    assert boxes.shape == (2, 75, 4) and scores.shape == (2, 75, 4)
    xr = x.reshape(2, 3, 9, 5, 5)
    sig = torch.sigmoid
    for (i, j, k, l) in [(0, 0, 0, 0), (1, 2, 3, 4), (0, 1, 4, 2)]:
        conf = float(sig(xr[i, j, 4, k, l]))
        idx = j * 25 + k * 5 + l
        if conf < 0.3:
            assert float(boxes[i, idx].abs().sum()) == 0 and float(scores[i, idx].abs().sum()) == 0
            continue
        ih, iw = float(img_size[i, 0]), float(img_size[i, 1])
        cx = (l + float(sig(xr[i, j, 0, k, l])) * 1.2 - 0.1) * iw / 5
        cy = (k + float(sig(xr[i, j, 1, k, l])) * 1.2 - 0.1) * ih / 5
        bw = math.exp(float(xr[i, j, 2, k, l])) * anchors[2 * j] * iw / 160
        bh = math.exp(float(xr[i, j, 3, k, l])) * anchors[2 * j + 1] * ih / 160
        want = [max(cx - bw / 2, 0), max(cy - bh / 2, 0), min(cx + bw / 2, iw - 1), min(cy + bh / 2, ih - 1)]
        np.testing.assert_allclose(boxes[i, idx].numpy(), want, rtol=1e-9)
        np.testing.assert_allclose(scores[i, idx].numpy(), conf * sig(xr[i, j, 5:, k, l]).numpy(), rtol=1e-9)
    xi = torch.randn(2, 3 + 3 * (5 + class_num), 5, 5, dtype=torch.float64)
    bi, si = O.yolo_box(xi, img_size, anchors, class_num, 0.0, 32, iou_aware=True, iou_aware_factor=0.4)
    conf = sig(xi[:, 3:].reshape(2, 3, 9, 5, 5)[:, :, 4]).pow(0.6) * sig(xi[:, :3]).pow(0.4)
    torch.testing.assert_close(si.reshape(2, 3, 5, 5, 4)[..., 0], conf * sig(xi[:, 3:].reshape(2, 3, 9, 5, 5)[:, :, 5]))
    # loss: every branch against the loop restatement - ignored cells, padded boxes, gt scores, a gt owned by another head,
    # two gts in one cell, label smoothing on and off
    anchors9 = [10, 13, 16, 30, 33, 23, 30, 61, 62, 45, 59, 119]
    gt_box = torch.tensor([[[0.31, 0.42, 0.10, 0.14], [0.33, 0.44, 0.16, 0.30], [0.7, 0.2, 0.35, 0.6], [0.0, 0.0, 0.0, 0.0]],
                           [[0.55, 0.51, 0.09, 0.2], [0.9, 0.9, 0.2, 0.15], [0.1, 0.8, 0.3, 0.25], [0.5, 0.5, 0.0, 0.1]]], dtype=torch.float64)
    gt_label = torch.tensor([[1, 0, 3, 0], [2, 2, 1, 0]], dtype=torch.int32)
    gt_score = torch.tensor([[1.0, 0.7, 0.9, 0.3], [0.5, 1.0, 0.8, 1.0]], dtype=torch.float64)
    for mask, smooth, score, sxy in (([0, 1, 2], True, None, 1.0), ([3, 4, 5], False, gt_score, 1.0), ([0, 1, 2], False, gt_score, 1.05)):
        xl = (torch.randn(2, 3 * (5 + class_num), 5, 5, dtype=torch.float64) * 0.8).requires_grad_(True)
        loss = O.yolo_loss(xl, gt_box, gt_label, anchors9, mask, class_num, 0.5, 32, gt_score=score, use_label_smooth=smooth, scale_x_y=sxy)
        want = _yolo_loss_loops(xl.detach().numpy(), gt_box.numpy(), gt_label.numpy(), (score if score is not None else torch.ones(2, 4)).numpy(),
                                anchors9, mask, class_num, 0.5, 32, smooth, sxy)
        np.testing.assert_allclose(loss.detach().numpy(), want, rtol=1e-9)
        loss.sum().backward()
        assert xl.grad is not None and torch.isfinite(xl.grad).all() and float(xl.grad.abs().sum()) > 0


def test_file_ops_and_conv_block(tmp_path):
    from PIL import Image
    arr = (np.random.RandomState(0).rand(12, 16, 3) * 255).astype("uint8")
    path = str(tmp_path / "a.jpg")
    Image.fromarray(arr).save(path, quality=95)
    raw = O.read_file(path)
    assert raw.dtype == torch.uint8 and raw.dim() == 1 and raw.numel() == len(open(path, "rb").read())
    img = O.decode_jpeg(raw)
    assert img.shape == (3, 12, 16) and img.dtype == torch.uint8
    np.testing.assert_array_equal(img.permute(1, 2, 0).numpy(), np.asarray(Image.open(path)))
    assert O.decode_jpeg(raw, mode="gray").shape == (1, 12, 16) and O.decode_jpeg(raw, mode="rgb").shape == (3, 12, 16)
    with pytest.raises(ValueError):
        O.decode_jpeg(raw, mode="cmyk")
    block = O.ConvNormActivation(3, 8, 3, stride=2)
    assert [type(m).__name__ for m in block] == ["Conv2D", "BatchNorm2D", "ReLU"] and block[0].bias is None
    assert block(torch.randn(2, 3, 9, 9)).shape == (2, 8, 5, 5)
    plain = O.ConvNormActivation(3, 8, 1, norm_layer=None, activation_layer=None)
    assert len(plain) == 1 and plain[0].bias is not None
    act = O.ConvNormActivation(3, 8, 3, dilation=2, activation_layer=paddle.nn.Hardswish)
    assert act(torch.randn(1, 3, 9, 9)).shape == (1, 8, 9, 9) and type(act[2]).__name__ == "Hardswish"
