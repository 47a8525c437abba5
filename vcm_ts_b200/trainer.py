"""The reference's multi-GPU training step for the P-frame model, on the drop-in `DMC`
(core/model/dcvc_hem.py:254-322 `forward_single_multi`, core/engine/train_multi.py:174-191 model wrap,
:203-240 `single_step`, trainer_multi.py:16-39 NCCL init, core/solver/optimizer.py:13 AdamW(0.9, 0.99)).

Pure data parallelism: every rank trains on its own crops, and the ONLY collective is the gradient
all-reduce of torch's DistributedDataParallel over NCCL / NVLink (bf16-compressed buckets); the forward and
backward of every convolution and warp are the device kernels behind vcm_ts_b200/training.py.  The
perceptual (detector-feature) loss of the reference is torchvision / ultralytics code outside the hot path
and is left to the caller (`perceptual_fn`).
"""
import math

import torch
from torch import nn

RATE_KEYS = {"none": [], "me": ["bpp_mv_y", "bpp_mv_z"], "rec": ["bpp_y", "bpp_z"],
             "all": ["bpp_mv_y", "bpp_mv_z", "bpp_y", "bpp_z"]}
DIST_KEYS = {"me": "me_mse", "rec": "mse"}


class DCVCTrainModule(nn.Module):
    """core/model/dcvc_hem.py `DCVC_HEM` reduced to the single-stage P-frame objective:
    loss = rate + lambda * (dist * dist_lambda + perceptual * pl_lambda), one lambda per anchor."""

    def __init__(self, dmc, lambdas, dist_lambda=1.0, pl_lambda=0.0, perceptual_fn=None):
        super().__init__()
        self.dmc = dmc
        self.register_buffer("lambdas", torch.tensor(list(lambdas), dtype=torch.float32))
        self.dist_lambda, self.pl_lambda, self.perceptual_fn = float(dist_lambda), float(pl_lambda), perceptual_fn

    def forward(self, x, target, dpb, loss_dist_key="mse", loss_rate_keys=RATE_KEYS["all"]):
        n = x.shape[0]
        assert self.lambdas.shape[0] == n, "batch = number of anchors (core/model/dcvc_hem.py:281)"
        lambdas = self.lambdas if len(loss_rate_keys) else torch.ones_like(self.lambdas)
        out = self.dmc.forward_one_frame(x, dpb, self.dmc.mv_y_q_scale, self.dmc.y_q_scale)
        new_dpb = {k: out["dpb"][k].detach() for k in dpb}                  # :290-291
        rate = torch.zeros_like(self.lambdas)
        for k in loss_rate_keys:
            rate = rate + out[k]
        dist = out[loss_dist_key]
        p_dist = torch.zeros_like(self.lambdas)
        if self.perceptual_fn is not None:
            p_dist = self.perceptual_fn(target, out["dpb"]["ref_frame"])
        loss = rate + lambdas * (dist * self.dist_lambda + p_dist * self.pl_lambda)
        return {"rate": rate, "dist": dist, "p_dist": p_dist, "loss": loss, "loss_to_opt": torch.mean(loss),
                "dpb": new_dpb}


def init_distributed(backend="nccl"):
    """trainer_multi.py:16-39: env:// rendezvous, one process per GPU."""
    import datetime
    import os
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    if not dist.is_initialized():
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device(f"cuda:{local}")
        dist.init_process_group(backend=backend, init_method="env://", timeout=datetime.timedelta(hours=5),
                                world_size=world, rank=rank, **kw)
    return rank, world, local


class Trainer:
    """init_model + single_step of core/engine/train_multi.py for the `single_multi` method."""

    def __init__(self, dmc, lambdas, lr=1e-4, world_size=1, device_ids=None, bf16_grad_buckets=True,
                 dist_lambda=1.0, pl_lambda=0.0, perceptual_fn=None):
        self.module = DCVCTrainModule(dmc, lambdas, dist_lambda, pl_lambda, perceptual_fn)
        self.module.to(next(dmc.parameters()).device)
        self.model = self.module
        if world_size > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            self.model = DDP(self.module, device_ids=device_ids, find_unused_parameters=True)   # :177
            if bf16_grad_buckets and device_ids is not None:
                from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
                self.model.register_comm_hook(None, default_hooks.bf16_compress_hook)
        # core/solver/optimizer.py:13, lr scaled by sqrt(world) (train_multi.py:160)
        self.optimizer = torch.optim.AdamW(self.module.parameters(), lr=lr * math.sqrt(world_size),
                                           betas=(0.9, 0.99))
        self.model.train()

    def single_step(self, input, target, t_i, p_frames=1, dist="rec", rate="all", dpb=None, sync=True):
        """One window of a [N,T,C,H,W] batch (train_multi.py:203-240): p_frames forward / backward /
        optimizer steps starting from frame t_i as the reference picture.  Returns the list of per-step
        results.  sync=False skips the gradient all-reduce (DDP no_sync; used to time the collective)."""
        import contextlib
        if dpb is None:
            dpb = {"ref_frame": input[:, t_i], "ref_feature": None, "ref_y": None, "ref_mv_y": None}
        results = []
        for p_idx in range(p_frames):
            self.optimizer.zero_grad()
            ctx = contextlib.nullcontext() if (sync or self.model is self.module) else self.model.no_sync()
            with ctx:
                res = self.model(input[:, t_i + 1 + p_idx], target[:, t_i + 1 + p_idx], dpb,
                                 DIST_KEYS[dist], RATE_KEYS[rate])
                res["loss_to_opt"].backward()
            self.optimizer.step()
            dpb = res["dpb"]
            results.append(res)
        return results
