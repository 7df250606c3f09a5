"""Training-state checkpoints interchangeable with the reference.

The reference snapshots a run with ``torch.save((gaussians.capture(), iteration), path)`` (train.py) and resumes with
``gaussians.restore(model_args, training_args)``; ``GaussianModel.capture`` (scene/gaussian_model.py:106-126) is an
18-tuple whose eleventh entry is ``optimizer.state_dict()``.  ``capture`` / ``restore`` here write and read exactly that
tuple, with ``FusedAdam`` (one launch per step, same ``state_dict`` layout as torch.optim.Adam) in the optimiser's
place, so a run started on the reference continues on libghr and the other way round.  ``training_setup`` builds the
reference's parameter groups (:250-287: names, order, learning rates).

Tensors stay where they are (CUDA); nothing here computes on the CPU.
"""
import torch

from .optim import FusedAdam

# order of GaussianModel.capture(), scene/gaussian_model.py:106-126
FIELDS = ("active_sh_degree", "xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity", "max_radii2D",
          "xyz_gradient_accum", "denom", "optimizer_state", "spatial_lr_scale", "pose_decoder",
          "lweight_offset_decoder", "normal", "albedo", "roughness")


def training_setup(state, training_args, motion_offset_flag=True):
    """GaussianModel.training_setup (scene/gaussian_model.py:250-287): the reference's parameter groups, in its order and
    with its learning rates, on FusedAdam(lr=0.0, eps=1e-15).  `state`: dict with the FIELDS above (parameters are leaf
    tensors).  Also (re)creates xyz_gradient_accum / denom like the reference does."""
    P = state["xyz"].shape[0]
    dev = state["xyz"].device
    state["percent_dense"] = training_args.percent_dense
    state["xyz_gradient_accum"] = torch.zeros((P, 1), device=dev)
    state["denom"] = torch.zeros((P, 1), device=dev)
    s = state["spatial_lr_scale"]
    l = [
        {"params": [state["xyz"]], "lr": training_args.position_lr_init * s, "name": "xyz"},
        {"params": [state["features_dc"]], "lr": training_args.feature_lr, "name": "f_dc"},
        {"params": [state["features_rest"]], "lr": training_args.feature_lr / 20.0, "name": "f_rest"},
        {"params": [state["opacity"]], "lr": training_args.opacity_lr, "name": "opacity"},
        {"params": [state["scaling"]], "lr": training_args.scaling_lr, "name": "scaling"},
        {"params": [state["rotation"]], "lr": training_args.rotation_lr, "name": "rotation"},
    ]
    if motion_offset_flag:
        state["normal"].requires_grad_(False)
        l += [
            {"params": list(state["pose_decoder"].parameters()), "lr": training_args.pose_refine_lr, "name": "pose_decoder"},
            {"params": list(state["lweight_offset_decoder"].parameters()), "lr": training_args.lbs_offset_lr,
             "name": "lweight_offset_decoder"},
            {"params": [state["normal"]], "lr": training_args.normal_lr, "name": "normal"},
            {"params": [state["albedo"]], "lr": training_args.opacity_lr, "name": "albedo"},
            {"params": [state["roughness"]], "lr": training_args.opacity_lr, "name": "roughness"},
        ]
    return FusedAdam(l, lr=0.0, eps=1e-15)


def capture(state, optimizer):
    """-> the tuple GaussianModel.capture() returns (scene/gaussian_model.py:106-126)."""
    out = []
    for f in FIELDS:
        out.append(optimizer.state_dict() if f == "optimizer_state" else state[f])
    return tuple(out)


def restore(model_args, training_args, motion_offset_flag=True):
    """GaussianModel.restore (scene/gaussian_model.py:128-150): -> (state dict, FusedAdam with the loaded state)."""
    if len(model_args) != len(FIELDS):
        raise ValueError(f"expected the {len(FIELDS)}-tuple of GaussianModel.capture(), got {len(model_args)} entries")
    state = dict(zip(FIELDS, model_args))
    opt_dict = state.pop("optimizer_state")
    xyz_gradient_accum, denom = state["xyz_gradient_accum"], state["denom"]
    optimizer = training_setup(state, training_args, motion_offset_flag)
    state["xyz_gradient_accum"], state["denom"] = xyz_gradient_accum, denom
    optimizer.load_state_dict(opt_dict)
    return state, optimizer
