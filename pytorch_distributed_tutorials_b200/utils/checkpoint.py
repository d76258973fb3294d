"""Checkpoint save / resume in the reference's format (reference resnet/main.py:84-85 load with
``map_location={"cuda:0": "cuda:<local_rank>"}``, :112 ``torch.save(ddp_model.state_dict(), path)``):
a plain ``torch.save`` of the wrapper's state_dict — ``module.``-prefixed torchvision keys, fp32 tensors plus
int64 ``num_batches_tracked`` — nothing else.  Extras the reference lacks are opt-in: the directory is created,
and optimizer / epoch state can ride along in a side file without changing the main file's format.
"""
import os

import torch


def save_checkpoint(model, path, optimizer=None, epoch=None):
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    # clone out of the flat buckets so the file holds exactly one tensor per key (not one giant shared storage)
    sd = {k: v.detach().clone().contiguous() for k, v in model.state_dict().items()}
    tmp = path + ".tmp"
    torch.save(sd, tmp)
    os.replace(tmp, path)
    if optimizer is not None or epoch is not None:
        extra = {"epoch": epoch, "optimizer": optimizer.state_dict() if optimizer is not None else None}
        torch.save(extra, path + ".extra")


def load_checkpoint(model, path, local_rank=None, strict=True, strip_prefix=False, optimizer=None):
    """Loads on every rank (as the reference does). ``strip_prefix`` lets a bare model read a wrapper's file."""
    map_location = None
    if local_rank is not None and torch.cuda.is_available():
        map_location = {"cuda:0": "cuda:%d" % local_rank}
    sd = torch.load(path, map_location=map_location, weights_only=True)
    if strip_prefix:
        sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}
    out = model.load_state_dict(sd, strict=strict)
    epoch = None
    if optimizer is not None and os.path.exists(path + ".extra"):
        extra = torch.load(path + ".extra", map_location=map_location, weights_only=False)
        if extra.get("optimizer") is not None:
            optimizer.load_state_dict(extra["optimizer"])
        epoch = extra.get("epoch")
    return out, epoch
