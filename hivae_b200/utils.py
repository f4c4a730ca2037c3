"""Checkpoint / training helpers with the reference's names and behaviour (model/utils.py): the on-disk format is
`checkpoint-<step>/model.safetensors` with the state-dict keys of SURVEY 8b, so reference checkpoints load unchanged into
the hivae_b200 modules (safetensors.torch.load_model, amd_inference.py:108-109)."""
import os
import re

from safetensors import safe_open


def find_latest_checkpoint(checkpoint_dir):
    """model/utils.py:217-237: <dir>/checkpoint-<largest step>/model.safetensors"""
    max_step, latest = -1, None
    for name in os.listdir(checkpoint_dir):
        m = re.match(r"checkpoint-(\d+)$", name)
        if m and int(m.group(1)) > max_step:
            max_step, latest = int(m.group(1)), os.path.join(checkpoint_dir, name)
    if latest is None:
        raise ValueError(f"No valid checkpoint found in {checkpoint_dir}")
    return os.path.join(latest, "model.safetensors")


def model_load_pretrain(model, path, not_load_keyword="decoder", strict=False):
    """model/utils.py:178-185: load every tensor whose key does not contain `not_load_keyword`"""
    tensors = {}
    with safe_open(path, framework="pt") as f:
        for k in f.keys():
            if not_load_keyword not in k:
                tensors[k] = f.get_tensor(k)
    return model.load_state_dict(tensors, strict=strict)


def frozen_model(model, frozen_name):
    """model/utils.py:172-176 (`--frozen_name camera_motion_encoder`, scripts/amd_new.sh:66)"""
    for name, module in model.named_modules():
        if frozen_name in name:
            for p in module.parameters():
                p.requires_grad_(False)


def print_param_num(model):
    """model/utils.py:103-114"""
    total = sum(p.numel() for p in model.parameters())
    trainable = sum(p.numel() for p in model.parameters() if p.requires_grad)
    print(f"total params: {total / 1e6:.2f} M, trainable: {trainable / 1e6:.2f} M")
    return total, trainable
