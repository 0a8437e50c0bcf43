"""One training step of the codec (BASELINE config 4) -- the step of /root/reference/train.py:195-234 and nothing else
(no loop, checkpoints, dataloader or validation: out of scope).

Data-parallel over one process per GPU: the model is replicated, every rank steps on its own blocks, and the gradients
are averaged with ONE bucketed NCCL all-reduce over NVLink before clipping (the reference has no distributed code at
all, train.py:56-57).  The `.quantiles` parameters have their own optimiser and aux-loss backward
(train.py:64-75,230-234); their 192x3 gradients ride in a second tiny all-reduce.
"""
import torch
import torch.distributed as dist

from . import MinkowskiEngine as ME
from .model.loss import Loss, default_loss_config


def scale_q_vals(q_vals, mode="quadratic", a_G=1.0, b_G=0.0, a_A=1.0, b_A=0.0):
    """q in [0,1]^2 -> Lambda map (data/q_func.py:49-71)."""
    lam = q_vals.clone()
    if mode == "exponential":
        lam[:, 0] = 2 ** (lam[:, 0] * a_G) + b_G
        lam[:, 1] = 2 ** (lam[:, 1] * a_A) + b_A
    elif mode == "quadratic":
        lam[:, 0] = lam[:, 0] ** 2 * a_G + b_G
        lam[:, 1] = lam[:, 1] ** 2 * a_A + b_A
    else:
        raise ValueError("Unknown mapping mode")
    return lam


def allreduce_mean_(params, world, bucket_bytes=64 << 20):
    """Average the gradients of `params` across ranks with flat fp32 buckets (one all-reduce for the ~72 MB model)."""
    if world <= 1:
        return 0
    grads = [p.grad for p in params if p.grad is not None]
    n_calls, i = 0, 0
    while i < len(grads):
        j, size = i, 0
        while j < len(grads) and (j == i or size + grads[j].numel() * 4 <= bucket_bytes):
            size += grads[j].numel() * 4
            j += 1
        flat = torch.cat([g.reshape(-1) for g in grads[i:j]])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(world)
        off = 0
        for g in grads[i:j]:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        n_calls += 1
        i = j
    return n_calls


class GradBuckets:
    """Bucketed gradient all-reduce launched DURING backward (overlap): every parameter gets a post-accumulate-grad hook; buckets
    are filled in the order gradients become ready (reverse of the forward order, as autograd produces them), and the moment the
    last gradient of a bucket has arrived the bucket is flattened and its asynchronous NCCL all-reduce is enqueued on the
    process group's own stream, where it runs under the remaining backward kernels.  `finish()` waits for the handles, divides
    by the world size and scatters the averages back into the .grad tensors.  Bucket membership is fixed at construction
    (parameter order), so every rank issues the same collectives in the same order."""

    def __init__(self, params, world, bucket_bytes=25 << 20):
        self.world = world
        self.params = [p for p in params if p.requires_grad]
        self.buckets, cur, size = [], [], 0
        for p in reversed(self.params):                    # backward order
            cur.append(p)
            size += p.numel() * 4
            if size >= bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self.bucket_of = {id(p): b for b, ps in enumerate(self.buckets) for p in ps}
        self.pending = [0] * len(self.buckets)
        self.inflight = []
        self.hooks = [p.register_post_accumulate_grad_hook(self._ready) for p in self.params] if world > 1 else []
        self.exposed_ms = 0.0

    def start(self):
        self.pending = [len(b) for b in self.buckets]
        self.inflight = []

    def _launch(self, b):
        grads = [p.grad for p in self.buckets[b] if p.grad is not None]
        if not grads:
            return
        flat = torch.cat([g.reshape(-1) for g in grads])
        self.inflight.append((dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=True), flat, grads))

    def _ready(self, p):
        b = self.bucket_of[id(p)]
        self.pending[b] -= 1
        if self.pending[b] == 0:
            self._launch(b)

    def finish(self):
        """Buckets whose parameters got no gradient this step never fired: launch them (with what they have), then wait."""
        if self.world <= 1:
            return 0
        for b, left in enumerate(self.pending):
            if left > 0:
                self.pending[b] = 0
                self._launch(b)
        for work, flat, grads in self.inflight:
            work.wait()
            flat.div_(self.world)
            off = 0
            for g in grads:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()
        n = len(self.inflight)
        self.inflight = []
        return n


class Trainer:
    """Optimisers + step.  Hyper-parameters are the `CVPR_*.yaml` ones (model lr 1e-4, bottleneck lr 1e-3, clip 1.0)."""

    def __init__(self, model, loss_config=None, model_lr=1e-4, bottleneck_lr=1e-3, clip_grad_norm=1.0, world=1, q_mode="quadratic",
                 q_params=None, overlap=True):
        self.model, self.world, self.clip = model, world, clip_grad_norm
        self.model_params = [p for n, p in model.named_parameters() if not n.endswith(".quantiles")]
        self.bottleneck_params = [p for n, p in model.named_parameters() if n.endswith(".quantiles")]
        self.model_optimizer = torch.optim.Adam(self.model_params, lr=model_lr)
        self.bottleneck_optimizer = torch.optim.Adam(self.bottleneck_params, lr=bottleneck_lr)
        self.loss = Loss(loss_config or default_loss_config())
        self.q_mode, self.q_params = q_mode, q_params or {}
        # gradient all-reduce overlapped with backward (hooks); overlap=False keeps the single flat all-reduce after backward
        self.buckets = GradBuckets(self.model_params, world) if (overlap and world > 1) else None

    def forward_backward(self, coords, feats, q):
        """Forward, losses, backward and the gradient all-reduce of the model parameters (their .grad then hold the mean over
        the ranks).  Returns (loss, parts)."""
        self.model.train()
        self.model_optimizer.zero_grad(set_to_none=True)
        self.bottleneck_optimizer.zero_grad(set_to_none=True)
        coords, feats = ME.utils.sparse_quantize(coordinates=coords, features=feats, quantization_size=1.0)
        x = ME.SparseTensor(features=feats, coordinates=coords, device=feats.device)
        lam = scale_q_vals(q, self.q_mode, **self.q_params)
        out = self.model(x, q, lam)
        loss, parts = self.loss(x, out)
        if self.buckets is not None:
            self.buckets.start()
            loss.backward()                                   # bucket all-reduces are enqueued from the hooks while this runs
            self.buckets.finish()
        else:
            loss.backward()
            allreduce_mean_(self.model_params, self.world)
        return loss, parts

    def step(self, coords, feats, q):
        """coords (N,4) int (batch,x,y,z), feats (N,3) rgb in [0,1], q (B,2).  Returns (loss, aux_loss, parts)."""
        loss, parts = self.forward_backward(coords, feats, q)
        torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.clip)
        self.model_optimizer.step()
        aux = self.model.aux_loss()
        aux.backward()
        allreduce_mean_(self.bottleneck_params, self.world)
        self.bottleneck_optimizer.step()
        return loss.detach(), aux.detach(), {k: v.detach() for k, v in parts.items()}
