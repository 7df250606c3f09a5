"""The LBS-weight offset network of the reference (nets/mlp_delta_weight_lbs.py:5-33 ``LBSOffsetDecoder``) on libghr.so.

``LBSOffsetDecoder`` here is state_dict-compatible with the reference's module (``bw_linears.{0..3}.{weight,bias}``,
``bw_fc.{weight,bias}``; Conv1d weights ``[out, in, 1]``) and is called the same way
(``decoder(means3D[None].detach())`` -> ``[1, 24, P]``, gaussian_renderer/__init__.py:105-106).  Its forward is ONE
tcgen05 / TMEM kernel (``ghr_lbs_mlp_forward``, csrc/ghr_mlp.cu: TF32 operands, fp32 accumulation -- the precision of the
reference's own Conv1d layers under torch's default ``cudnn.allow_tf32``); the weights are repacked for the tensor core
whenever they change (``ghr_lbs_mlp_pack``, one small launch).  The backward (gradients of the network's parameters; the
reference detaches the positions) recomputes the layers with torch's own convolution kernels: a library path, like the
reference's.  There is no CPU fallback.
"""
import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from ._lib import check, lib
from .lbs import _stream

N_FREQS = 10
EMB = 3 + 3 * 2 * N_FREQS  # 63


def xyz_embedder(pts):
    """get_embedder(10) (mlp_delta_weight_lbs.py:36-83): [x, sin(2^f x), cos(2^f x)] for f = 0..9 -> [..., 63]."""
    out = [pts]
    for f in range(N_FREQS):
        out += [torch.sin(pts * 2.0 ** f), torch.cos(pts * 2.0 ** f)]
    return torch.cat(out, -1)


class _MlpFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, module, pts, *params):
        if not pts.is_cuda:
            raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
        x = pts.detach().reshape(-1, 3)
        x = (x if x.dtype == torch.float32 else x.float()).contiguous()
        P = x.shape[0]
        logits = torch.empty(P, 24, dtype=torch.float32, device=x.device)
        blob = module._packed(params)
        if P:
            with _lib.on_device(x.device):
                check(lib().ghr_lbs_mlp_forward(P, x.data_ptr(), blob.data_ptr(), logits.data_ptr(), _stream()),
                      "ghr_lbs_mlp_forward")
        ctx.module, ctx.pts = module, pts
        ctx.save_for_backward(*params)
        return logits

    @staticmethod
    def backward(ctx, g):
        params = ctx.saved_tensors
        need = ctx.needs_input_grad
        pts = ctx.pts.detach().requires_grad_(need[1])
        with torch.enable_grad():
            ps = [p.detach().requires_grad_(n) for p, n in zip(params, need[2:])]
            out = ctx.module.torch_forward(pts, ps)           # [1, 24, P]
            wanted = [t for t in [pts] + ps if t.requires_grad]
            grads = torch.autograd.grad(out, wanted, g.t()[None]) if wanted else []
        it = iter(grads)
        return (None,) + tuple(next(it) if t.requires_grad else None for t in [pts] + ps)


class LBSOffsetDecoder(nn.Module):
    def __init__(self, total_bones=24):
        super().__init__()
        assert total_bones == 24, "the kernel is built for SMPL's 24 bones"
        self.total_bones, self.skips = total_bones, [2]
        Wd = 128
        self.bw_linears = nn.ModuleList([nn.Conv1d(EMB, Wd, 1), nn.Conv1d(Wd, Wd, 1), nn.Conv1d(Wd, Wd, 1),
                                         nn.Conv1d(Wd + EMB, Wd, 1)])
        self.bw_fc = nn.Conv1d(Wd, total_bones, 1)
        self._blob, self._blob_key = None, None

    def _params(self):
        return [l.weight for l in self.bw_linears] + [self.bw_fc.weight] + [l.bias for l in self.bw_linears] + [self.bw_fc.bias]

    def _packed(self, params):
        """The tensor-core blob of the current parameter values (repacked when any of them changed)."""
        key = tuple((p.data_ptr(), p._version) for p in params)
        if key != self._blob_key or self._blob is None:
            dev = params[0].device
            l = lib()
            if self._blob is None or self._blob.device != dev:
                self._blob = torch.empty(l.ghr_lbs_mlp_blob_bytes(), dtype=torch.uint8, device=dev)
            tensors = [p.detach().float().contiguous() for p in params]
            w = (C.c_void_p * 5)(*[t.data_ptr() for t in tensors[:5]])
            b = (C.c_void_p * 5)(*[t.data_ptr() for t in tensors[5:]])
            with _lib.on_device(dev):
                check(l.ghr_lbs_mlp_pack(w, b, self._blob.data_ptr(), _stream()), "ghr_lbs_mlp_pack")
            if not torch.cuda.is_current_stream_capturing():
                self._blob_key = key
        return self._blob

    def torch_forward(self, pts, params=None):
        """The reference's forward statement by statement (used for the backward and as the in-package cross-check)."""
        p = self._params() if params is None else params
        features = xyz_embedder(pts).permute(0, 2, 1)
        net = features
        for i in range(4):
            net = F.relu(F.conv1d(net, p[i], p[5 + i]))
            if i in self.skips:
                net = torch.cat((features, net), dim=1)
        return F.conv1d(net, p[4], p[9])

    def forward(self, pts):
        """pts [1, P, 3] -> [1, 24, P] (a view of the kernel's [P, 24] logits: the caller's permute(0, 2, 1) gives them
        back row-major, which is what the deform kernels read)."""
        assert pts.dim() == 3 and pts.shape[0] == 1 and pts.shape[2] == 3, "batch size 1, as the reference uses it"
        return _MlpFn.apply(self, pts, *self._params()).t()[None]


# ---- BodyPoseRefiner (nets/mlp_delta_body_pose.py:5-41) ---------------------------------------------------------------

def _rodrigues_1e5(rvec):
    """RodriguesModule.forward (mlp_delta_body_pose.py:164-193): theta = sqrt(1e-5 + |r|^2)."""
    theta = torch.sqrt(1e-5 + torch.sum(rvec ** 2, dim=1))
    r = rvec / theta[:, None]
    c, s = torch.cos(theta), torch.sin(theta)
    return torch.stack((
        r[:, 0] ** 2 + (1. - r[:, 0] ** 2) * c, r[:, 0] * r[:, 1] * (1. - c) - r[:, 2] * s, r[:, 0] * r[:, 2] * (1. - c) + r[:, 1] * s,
        r[:, 0] * r[:, 1] * (1. - c) + r[:, 2] * s, r[:, 1] ** 2 + (1. - r[:, 1] ** 2) * c, r[:, 1] * r[:, 2] * (1. - c) - r[:, 0] * s,
        r[:, 0] * r[:, 2] * (1. - c) - r[:, 1] * s, r[:, 1] * r[:, 2] * (1. - c) + r[:, 0] * s, r[:, 2] ** 2 + (1. - r[:, 2] ** 2) * c),
        dim=1).view(-1, 3, 3)


class _RefinerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, module, pose, *params):
        if not pose.is_cuda:
            raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
        x = pose.detach().reshape(-1).float().contiguous()
        assert x.numel() == 69, "one pose vector (23 x 3) per call, as render() uses it"
        n = len(params) // 2
        tensors = [p.detach().float().contiguous() for p in params]
        w = (C.c_void_p * n)(*[t.data_ptr() for t in tensors[:n]])
        b = (C.c_void_p * n)(*[t.data_ptr() for t in tensors[n:]])
        Rs = torch.empty(1, 23, 3, 3, dtype=torch.float32, device=x.device)
        with _lib.on_device(x.device):
            check(lib().ghr_pose_refiner_forward(module.mlp_width, n, w, b, x.data_ptr(), Rs.data_ptr(), _stream()),
                  "ghr_pose_refiner_forward")
        ctx.module, ctx.pose = module, pose
        ctx.save_for_backward(*params)
        return Rs

    @staticmethod
    def backward(ctx, g):
        params = ctx.saved_tensors
        need = ctx.needs_input_grad
        pose = ctx.pose.detach().requires_grad_(need[1])
        with torch.enable_grad():
            ps = [p.detach().requires_grad_(n) for p, n in zip(params, need[2:])]
            out = ctx.module.torch_forward(pose, ps)
            wanted = [t for t in [pose] + ps if t.requires_grad]
            grads = torch.autograd.grad(out, wanted, g) if wanted else []
        it = iter(grads)
        return (None,) + tuple(next(it) if t.requires_grad else None for t in [pose] + ps)


class BodyPoseRefiner(nn.Module):
    """state_dict-compatible with the reference's module (``block_mlps.{0,2,...}.{weight,bias}``); same constructor
    arguments, same call: ``refiner(poses[:, 3:])["Rs"]`` -> [1, 23, 3, 3].  Forward = one kernel
    (ghr_pose_refiner_forward); the backward recomputes through torch.  The reference's initialisation (Xavier with ReLU
    gain, last layer ~0 so that the corrections start at identity) is reproduced."""

    def __init__(self, total_bones=24, embedding_size=69, mlp_width=256, mlp_depth=4, **_):
        super().__init__()
        assert total_bones == 24 and embedding_size == 69, "built for SMPL's 23 non-root joints"
        import math
        self.total_bones, self.mlp_width = total_bones - 1, mlp_width
        layers = [nn.Linear(embedding_size, mlp_width), nn.ReLU()]
        for _i in range(mlp_depth - 1):
            layers += [nn.Linear(mlp_width, mlp_width), nn.ReLU()]
        layers += [nn.Linear(mlp_width, 3 * self.total_bones)]
        self.block_mlps = nn.Sequential(*layers)
        lin = [m for m in self.block_mlps if isinstance(m, nn.Linear)]
        for i, m in enumerate(lin):  # initseq: ReLU gain for layers followed by a ReLU, 1 for the last
            gain = math.sqrt(2.0) if i < len(lin) - 1 else 1.0
            std = gain * math.sqrt(2.0 / (m.in_features + m.out_features))
            m.weight.data.uniform_(-(std * math.sqrt(3.0)), std * math.sqrt(3.0))
            m.bias.data.zero_()
        lin[-1].weight.data.uniform_(-1e-5, 1e-5)
        lin[-1].bias.data.zero_()

    def _params(self):
        lin = [m for m in self.block_mlps if isinstance(m, nn.Linear)]
        return [m.weight for m in lin] + [m.bias for m in lin]

    def torch_forward(self, pose_input, params=None):
        p = self._params() if params is None else params
        n = len(p) // 2
        x = pose_input
        for i in range(n):
            x = F.linear(x, p[i], p[n + i])
            if i < n - 1:
                x = F.relu(x)
        return _rodrigues_1e5(x.view(-1, 3)).view(-1, self.total_bones, 3, 3)

    def forward(self, pose_input):
        return {"Rs": _RefinerFn.apply(self, pose_input, *self._params())}
