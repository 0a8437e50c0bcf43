"""CPU restatement of the reference's codec pipeline on top of the oracle.  TEST INFRASTRUCTURE and
bench.py's timed CPU baseline ("the reference's CPU MinkowskiEngine (OpenBLAS) path").

Follows, op for op:
  /root/reference/model/transforms.py:68-97   AnalysisTransform.forward      -> g_a
  /root/reference/model/transforms.py:170-282 SparseSynthesisTransform       -> g_s
  /root/reference/model/blocks.py:26-57       MinkowskiGDN                   -> _gdn
  /root/reference/model/entropy_models.py:344-490 compress / decompress     -> entropy_encode / entropy_decode
  /root/reference/model/model.py:94-250       UnifiedModel.compress/decompress -> compress / decompress
Sparse convolutions run through oracle/csrc/sparse_conv_ref.cpp (hash map -> kernel map -> per-offset
gather / OpenBLAS sgemm / scatter-add, all host cores); GDN, entropy-model and MLP arithmetic is
PyTorch-CPU eager exactly as the reference executes it.  Weights come from a state dict with the
reference's key names.  PARITY UNPINNED by the reference itself (oracle/__init__.py); pinned against the
reference's own model/*.py run over oracle/me_ref.py + cai_ref.py by tests/golden (oracle/make_golden.py).
"""
import numpy as np
import torch

from . import cai_ref, me_ref, native


def _np(t):
    return t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


class RefCodec:
    def __init__(self, state_dict, config):
        self.sd = {k: v.detach().cpu().float() if v.is_floating_point() else v.detach().cpu() for k, v in state_dict.items()}
        self.cfg = config
        em = config["entropy_model"]
        self.inverse_rescaling = em["inverse_rescaling"]
        self.quantization_offset = em["quantization_offset"]
        self.adaptive_BN = em.get("adaptive_BN", True)
        self.eb = cai_ref.EntropyBottleneck(em["C_hyper_bottleneck"])
        self.gc = cai_ref.GaussianConditional(None)
        self._load(self.eb, "entropy_model.entropy_bottleneck.")
        self.eb.update(force=True)
        self.gc.update_scale_table(cai_ref.get_scale_table(), force=True)
        self.beta_rp = cai_ref.NonNegativeParametrizer(minimum=1e-6)
        self.gamma_rp = cai_ref.NonNegativeParametrizer()

    def _load(self, module, prefix):
        sub = {k[len(prefix):]: v for k, v in self.sd.items() if k.startswith(prefix) and not k[len(prefix):].startswith("_")}
        module.load_state_dict(sub, strict=False)

    # -- layers ----------------------------------------------------------------------------------
    def _w(self, name):
        return self.sd[name + ".kernel"], self.sd.get(name + ".bias")

    def _conv(self, name, C, F, ts, ksize, stride=1, act=0):
        W, b = self._w(name)
        C_out = C if stride == 1 else native.stride_coords(C, ts * stride)
        nbr = native.kernel_map(C, C_out, ksize, ts, False)
        return C_out, native.conv(F, nbr, _np(W), None if b is None else _np(b), act), ts * stride

    def _gen(self, name, C, F, ts, ksize, stride=2, act=0):
        W, b = self._w(name)
        ts_out = ts // stride
        C_out = native.generate_coords(C, ksize, ts_out)
        nbr = native.kernel_map(C, C_out, ksize, ts_out, True)
        return C_out, native.conv(F, nbr, _np(W), None if b is None else _np(b), act), ts_out

    def _gdn(self, name, F, inverse):
        x = torch.from_numpy(F)
        beta = self.beta_rp(self.sd[name + ".beta"])
        gamma = self.gamma_rp(self.sd[name + ".gamma"])
        C = x.shape[1]
        norm = torch.nn.functional.conv1d(torch.abs(x.T.unsqueeze(0)), gamma.reshape(C, C, 1), beta)[0].T
        if not inverse:
            norm = 1.0 / norm
        return (x * norm).numpy()

    def _mlp(self, prefix, x, idxs, last=None):
        for j, i in enumerate(idxs):
            x = torch.nn.functional.linear(x, self.sd[f"{prefix}.{i}.weight"], self.sd[f"{prefix}.{i}.bias"])
            if j < len(idxs) - 1:
                x = torch.relu(x)
        return last(x) if last is not None else x

    # -- transforms --------------------------------------------------------------------------------
    def g_a(self, C, F):
        k = [[len(C)]]
        C, F, ts = self._conv("g_a.down_conv_1.0", C, F, 1, 5, 2)
        F = self._gdn("g_a.down_conv_1.1", F, False)
        k.append([len(C)])
        C, F, ts = self._conv("g_a.down_conv_2.0", C, F, ts, 5, 2)
        F = self._gdn("g_a.down_conv_2.1", F, False)
        k.append([len(C)])
        C, F, ts = self._conv("g_a.down_conv_3.0", C, F, ts, 5, 2)
        F = self._gdn("g_a.down_conv_3.1", F, False)
        C, F, ts = self._conv("g_a.down_conv_3.2", C, F, ts, 5, 1)
        k.reverse()
        return C, F, k

    @staticmethod
    def topk_mask(logits, k):
        """torch.topk on the logits of the (single) batch item (transforms.py:228-254)."""
        mask = np.zeros(len(logits), bool)
        _, top = torch.topk(torch.from_numpy(np.ascontiguousarray(logits)), int(min(k, len(logits))))
        mask[top.numpy()] = True
        return mask

    def g_s(self, C, F, k, return_logits=False):
        ts = 8
        logits_all = []
        C, F, ts = self._conv("g_s.up_1.0", C, F, ts, 5, 1)
        ups = (("g_s.up_1.1", "g_s.up_1.2", "g_s.predict_1"), ("g_s.up_2.0", "g_s.up_2.1", "g_s.predict_2"),
               ("g_s.up_3.0", "g_s.up_3.1", "g_s.predict_3"))
        for level, (gdn, gen, head) in enumerate(ups):
            F = self._gdn(gdn, F, True)
            C, F, ts = self._gen(gen, C, F, ts, 5, 2)
            _, h, _ = self._conv(head + ".0", C, F, ts, 3, 1, act=1)
            _, logit, _ = self._conv(head + ".2", C, h, ts, 3, 1)
            logits_all.append((C, logit))
            mask = self.topk_mask(logit[:, 0], k[level][0])
            C, F = C[mask], F[mask]
        _, F, _ = self._conv("g_s.color_conv.0", C, F, ts, 1, 1)
        return (C, F, logits_all) if return_logits else (C, F)

    # -- entropy model -----------------------------------------------------------------------------
    def h_a(self, C, F):
        C1, F1, ts = self._conv("entropy_model.h_a.0", C, F, 8, 3, 1, act=2)
        C2, F2, ts = self._conv("entropy_model.h_a.2", C1, F1, ts, 3, 2, act=2)
        C3, F3, ts = self._conv("entropy_model.h_a.4", C2, F2, ts, 3, 2)
        return C3, F3

    def h_s(self, C, F):
        C1, F1, ts = self._gen("entropy_model.h_s.0", C, F, 32, 2, 2, act=2)
        C2, F2, ts = self._gen("entropy_model.h_s.2", C1, F1, ts, 2, 2, act=2)
        C3, F3, ts = self._conv("entropy_model.h_s.4", C2, F2, ts, 3, 1)
        return C3, F3

    def _gains(self, q, n):
        Cb = self.cfg["entropy_model"]["C_bottleneck"]
        if not self.adaptive_BN:
            one = torch.ones(1, Cb, n)
            return one, one
        qt = torch.as_tensor(q, dtype=torch.float32).reshape(1, 2)
        scale = self._mlp("entropy_model.scale_nn", qt, (0, 2, 4), torch.nn.functional.softplus) + 0.0001
        scale = scale[torch.zeros(n, dtype=torch.long)].t().unsqueeze(0)
        if self.inverse_rescaling:
            rescale = torch.tensor(1.0) / scale
        else:
            r = torch.tensor(1.0) / self._mlp("entropy_model.rescale_nn", qt, (0, 2, 4), torch.nn.functional.softplus)
            rescale = r[torch.zeros(n, dtype=torch.long)].t().unsqueeze(0)
        return scale, rescale

    def _gaussian_params(self, Cz, z_hat, Cy):
        Cg, Fg = self.h_s(Cz, z_hat)
        look = me_ref._Lookup(Cg)
        rows = look(Cy)
        feats = np.where(rows[:, None] >= 0, Fg[np.maximum(rows, 0)], 0.0).astype(np.float32)   # on-grid lookup (ME6)
        g = torch.from_numpy(feats)
        scales_hat, means_hat = g.chunk(2, dim=1)
        return scales_hat.t().unsqueeze(0), means_hat.t().unsqueeze(0)

    def entropy_encode(self, Cy, Fy, q):
        """-> dict with sorted points, int32 symbols/indexes in bitstream order (channel-major)."""
        Cz, Fz = self.h_a(Cy, Fy)
        py, pz = me_ref.lexsort_rows(Cy), me_ref.lexsort_rows(Cz)
        Cy, Fy, Cz, Fz = Cy[py], Fy[py], Cz[pz], Fz[pz]
        z = torch.from_numpy(Fz).t().unsqueeze(0)
        med = self.eb._get_medians().detach().reshape(1, -1, 1)
        z_sym = torch.round(z - med).int()
        z_hat = (z_sym.float() + med)[0].t().contiguous().numpy()
        scales_hat, means_hat = self._gaussian_params(Cz, z_hat, Cy)
        scale, _ = self._gains(q, len(Cy))
        indexes = self.gc.build_indexes(scales_hat * scale)
        y = torch.from_numpy(Fy).t().unsqueeze(0)
        y_sym = torch.round(y * scale - means_hat * scale).int()
        return {"y_points": Cy, "z_points": Cz, "z_symbols": z_sym.numpy(), "y_symbols": y_sym.numpy(),
                "y_indexes": indexes.numpy()}

    def entropy_decode(self, Cy, Cz, z_sym, y_sym, q):
        Cy, Cz = Cy[me_ref.lexsort_rows(Cy)], Cz[me_ref.lexsort_rows(Cz)]
        med = self.eb._get_medians().detach().reshape(1, -1, 1)
        z_hat = (torch.from_numpy(z_sym).float() + med)[0].t().contiguous().numpy()
        scales_hat, means_hat = self._gaussian_params(Cz, z_hat, Cy)
        scale, rescale = self._gains(q, len(Cy))
        indexes = self.gc.build_indexes(scales_hat * scale)
        if self.quantization_offset:
            q_val = torch.from_numpy(y_sym).float()
            q_abs, signs = q_val.abs(), torch.sign(q_val)
            sd = self.gc.lower_bound_scale(scales_hat * scale)
            comb = torch.cat((scale.unsqueeze(3), sd.unsqueeze(3)), dim=3)
            off = -self._mlp("entropy_model.quant_nn", comb, (0, 2, 4)).squeeze(3)
            off[q_abs < 0.0001] = 0
            y_hat = signs * (q_abs + off) * rescale + means_hat
        else:
            y_hat = torch.from_numpy(y_sym).float() + means_hat * scale
        return Cy, y_hat[0].t().contiguous().numpy(), indexes.numpy()

    # -- block level ---------------------------------------------------------------------------------
    def encode_block(self, xyz, rgb, q):
        C = np.concatenate([np.zeros((len(xyz), 1), np.int32), np.floor(xyz).astype(np.int32)], axis=1)
        first = me_ref.unique_first(C)
        C, rgb = C[first], np.asarray(rgb, np.float32)[first]
        F = np.concatenate([np.ones((len(C), 1), np.float32), rgb], axis=1)
        Cy, Fy, k = self.g_a(C, F)
        enc = self.entropy_encode(Cy, Fy, q)
        enc["k"] = k
        return enc

    def decode_block(self, y_points, z_sym, y_sym, k, q):
        Cz = native.stride_coords(native.stride_coords(y_points, 16), 32)
        Cy, y_hat, _ = self.entropy_decode(y_points, Cz, z_sym, y_sym, q)
        C, F = self.g_s(Cy, y_hat, k)
        rgb = np.clip(np.round(F * 255), 0.0, 255.0) / 255
        return np.concatenate([C[:, 1:4].astype(np.float32), rgb.astype(np.float32)], axis=1)
