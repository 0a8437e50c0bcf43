"""Training losses -- mirror of /root/reference/loss.py:9-274 (Loss wrapper, BPPLoss, ColorLoss, Multiscale_FocalLoss,
ShepardsLoss).  Same config keys, same arithmetic.  What changes is the set-membership machinery: the reference builds
int64 keys and calls `torch.isin` (sort-based, loss.py:99-103,143-147; utils.py:672-703); here membership is a probe
of the other tensor's coordinate hash (`CoordSet.lookup`), and the Shepard channel-wise 9^3 convolution (loss.py:185-188)
is one kernel (`pcg_channelwise_conv`) instead of 729 Python-loop iterations.
"""
import math

import torch

from .. import MinkowskiEngine as ME


def overlapping_mask(a, b):
    """Rows of `a` whose coordinate also occurs in `b` (utils.py:672-703)."""
    return b.coord_set.lookup(a.C) >= 0


class BPPLoss:
    def __init__(self, config):
        self.weight, self.identifier, self.key = config["weight"], config["id"], config["key"]

    def __call__(self, gt, pred):
        likelihoods = pred["likelihoods"][self.key]
        n = gt.C.shape[0]
        loss = 0.0
        for lk in likelihoods:                      # iterates the leading (batch = 1) dimension, loss.py:77-79
            loss = loss + torch.log(lk).sum() / (-math.log(2) * n)
        return loss.mean() * self.weight


class ColorLoss:
    def __init__(self, config):
        self.identifier = config["id"]
        self.loss_func = torch.nn.L1Loss(reduction="none") if config["loss"] == "L1" else torch.nn.MSELoss(reduction="none")

    def __call__(self, gt, pred):
        prediction, q_map = pred["prediction"], pred["q_map"]
        mask = overlapping_mask(gt, prediction)
        if not bool(mask.any()):        # no predicted voxel on the ground truth (the reference's mean() of an empty tensor is NaN)
            return prediction.F.sum() * 0.0
        pred_colors = prediction.features_at_coordinates(gt.C[mask].float())
        batch = gt.C[mask, 0].long()
        return (self.loss_func(gt.F[mask], pred_colors) * q_map[batch, 1].unsqueeze(1)).mean()


class Multiscale_FocalLoss:
    def __init__(self, config):
        self.identifier, self.alpha, self.gamma = config["id"], config["alpha"], config["gamma"]

    def __call__(self, gt, pred):
        predictions, points = list(pred["occ_predictions"]), list(pred["points"])
        predictions.reverse()
        points.reverse()
        q_map = pred["q_map"]
        loss = 0.0
        for prediction, coords in zip(predictions, points):
            occupied = overlapping_mask(prediction, coords)
            p_z = torch.sigmoid(prediction.F[:, 0])
            pt_z = torch.clip(torch.where(occupied, p_z, 1 - p_z), 1e-2, 1)
            alpha_z = torch.where(occupied, self.alpha, 1 - self.alpha)
            focal = -alpha_z * (1 - pt_z) ** self.gamma * torch.log(pt_z)
            loss = loss + (focal * q_map[prediction.C[:, 0].long(), 0]).mean()
        return loss


class ShepardsLoss:
    """Inverse-distance interpolation of GT colours onto predicted-only voxels, then L1/L2 (loss.py:161-274)."""

    def __init__(self, config):
        self.identifier = config["id"]
        self.loss_func = torch.nn.L1Loss(reduction="none") if config["loss"] == "L1" else torch.nn.MSELoss(reduction="none")
        self.p, self.window_size = config["p"], config["window_size"]
        self.conv_sum = ME.MinkowskiChannelwiseConvolution(in_channels=4, kernel_size=self.window_size, stride=1, dimension=3)
        self.conv_sum.kernel = torch.nn.Parameter(self.create_window_3D(self.window_size), requires_grad=False)

    def create_window_3D(self, w):
        r = w // 2
        z, y, x = torch.meshgrid(torch.arange(w) - r, torch.arange(w) - r, torch.arange(w) - r, indexing="ij")
        d = torch.sqrt((x ** 2 + y ** 2 + z ** 2).float())
        window = 1 / (d ** self.p + 1e-5)
        window[d > r] = 0
        return window.view(-1, 1)

    def interpolate_gt_to_pred(self, gt, prediction):
        dev = gt.C.device
        on_gt = overlapping_mask(prediction, gt)
        combined = torch.cat([gt.C, prediction.C[~on_gt]])
        feats = torch.zeros((combined.shape[0], 4), device=dev)
        feats[: gt.C.shape[0], 0] = 1.0
        feats[: gt.C.shape[0], 1:] = gt.F
        comb = ME.SparseTensor(features=feats, coordinates=combined, device=dev)
        interp = self.conv_sum.to(dev)(comb)
        raw = interp.features_at_coordinates(prediction.C[~on_gt].float())
        out = torch.zeros((prediction.F.shape[0], 3), device=dev)
        out[on_gt] = gt.features_at_coordinates(prediction.C[on_gt].float())
        out[~on_gt] = raw[:, 1:] / raw[:, 0].unsqueeze(1)
        return ME.SparseTensor(out, _coord_set=prediction.coord_set)

    def __call__(self, gt, pred):
        prediction, q_map = pred["prediction"], pred["q_map"]
        with torch.no_grad():
            gt_on_pred = self.interpolate_gt_to_pred(gt, prediction)
            valid = (~torch.isnan(gt_on_pred.F) & ~torch.isinf(gt_on_pred.F)).all(dim=1)
        if not bool(valid.any()):       # every predicted voxel is farther than the window radius from the ground truth
            return prediction.F.sum() * 0.0
        batch = prediction.C[valid, 0].long()
        return (self.loss_func(gt_on_pred.F[valid], prediction.F[valid]) * q_map[batch, 1].unsqueeze(1)).mean()


class Loss:
    """loss.py:9-59: sum of the configured losses."""
    TYPES = {"BPPLoss": BPPLoss, "ColorLoss": ColorLoss, "ShepardsLoss": ShepardsLoss, "Multiscale_FocalLoss": Multiscale_FocalLoss}

    def __init__(self, config):
        self.losses = {}
        for name, setting in config.items():
            setting = dict(setting, id=name)
            if setting["type"] in self.TYPES:
                self.losses[name] = self.TYPES[setting["type"]](setting)

    def __call__(self, gt, pred):
        total, parts = 0, {}
        for loss in self.losses.values():
            item = loss(gt, pred)
            parts[loss.identifier] = item
            total = total + item
        return total, parts


def default_loss_config(shepard=False):
    """`loss:` section of configs/CVPR_inverse_scaling.yaml (and _shepard.yaml)."""
    color = {"type": "ShepardsLoss", "loss": "L2", "window_size": 9, "p": 8} if shepard else {"type": "ColorLoss", "loss": "L2"}
    return {"Multiscale_FocalLoss": {"type": "Multiscale_FocalLoss", "alpha": 0.5, "gamma": 2.0},
            "ColorLoss": color,
            "bpp-y": {"type": "BPPLoss", "key": "y", "weight": 1.0},
            "bpp-z": {"type": "BPPLoss", "key": "z", "weight": 1.0}}
