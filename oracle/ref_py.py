"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Imports the reference's OWN Python modules of the render hot path from baseline/_ref/ (unmodified copies made by
oracle/install_ref_py.sh in the dev container, where /root/reference exists; git-ignored, shipped to the GPU box):

    scene/gaussian_model.py      GaussianModel.coarse_deform_c2source (:768-872), get_covariance (:208-209),
                                 capture/restore (:106-150), get_transform_params_torch (:947-980)
    utils/sh_utils.py            eval_sh (:135-190)
    gaussian_renderer/__init__.py  render() (:53-295)
    baking.py                    bake_set (:136-309)
    nets/*.py                    LBSOffsetDecoder, BodyPoseRefiner

Used by bench.py's reference arm (so that its LBS / covariance / SH half is the reference's own code, not a
restatement) and by the drop-in tests.  Third-party packages the reference imports that are absent from this image
are replaced by the smallest possible stand-ins, each named here:

    knn_cuda.KNN        (KNN_CUDA 0.2, un-vendored)  -> exact brute-force nearest neighbours (k + 2 GEMM-form candidates, exact re-rank)
    plyfile, trimesh, imageio   (file I/O, unused on the path)  -> empty modules
    nvdiffrast.torch    (un-vendored)  -> `texture` is whatever the caller installs (see set_nvdiffrast_texture)
    gs_ir._C            (the fork's private extension, unused by bake_set)  -> empty module
    scene.Scene         (dataset readers: cv2/imageio/PIL chains)  -> None
    simple_knn._C / diff_gaussian_rasterization  -> whatever the caller installs in sys.modules BEFORE calling
                        load() (mygauhuman_b200.install() for the drop-in test; an oracle/_ref adapter for the
                        reference arm)
"""
import importlib
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PY = os.path.join(ROOT, "baseline", "_ref")

_loaded = None


def available():
    return os.path.exists(os.path.join(REF_PY, "scene", "gaussian_model.py"))


def _stub(name, **attrs):
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        sys.modules[name] = m
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


class BruteForceKNN:
    """knn_cuda.KNN(k, transpose_mode=True)(ref[1,V,3], query[1,P,3]) -> (dist[1,P,k], idx[1,P,k]), exact."""

    def __init__(self, k, transpose_mode=True):
        self.k = k

    def __call__(self, ref, query):
        import torch
        # Exact, and as cheap as torch allows (the reference arm must not be charged for its missing dependency; measured
        # on B200 at 50 000 x 6 890: 2.4 ms, where torch.topk on the distance matrix costs 4.3 ms and an all-pairs exact
        # cdist 410 ms): candidates = the k + 2 smallest matmul-form distances (|a|^2 + |b|^2 - 2ab, one GEMM; off by
        # ~1e-6 m^2, which by itself picks another vertex for ~0.06 % of the Gaussians), found by k + 2 row minima with
        # the previous winners masked out, then re-ranked by exact differences like KNN_CUDA's kernel.
        n_cand = min(ref.shape[1], self.k + 2)
        out_d, out_i = [], []
        for s in range(0, query.shape[1], 65536):  # bounded distance matrix
            q = query[:, s:s + 65536]
            d = torch.cdist(q, ref)[0]                                                          # [p, V]
            cand = []
            for _ in range(n_cand):
                idx = d.argmin(-1, keepdim=True)
                cand.append(idx)
                d.scatter_(1, idx, float("inf"))
            cand = torch.cat(cand, 1)[None]                                                     # [1, p, n_cand]
            pts = ref[0][cand[0]]                                                               # [p, n_cand, 3]
            exact = (pts - q[0][:, None, :]).square().sum(-1).sqrt()[None]                      # [1, p, n_cand]
            dist, pick = torch.topk(exact, self.k, dim=-1, largest=False)
            out_d.append(dist); out_i.append(torch.gather(cand, -1, pick))
        return torch.cat(out_d, 1), torch.cat(out_i, 1)


def set_nvdiffrast_texture(fn):
    """Install `fn(tex, uv, filter_mode=..., boundary_mode=...)` as nvdiffrast.torch.texture for bake_set."""
    _stub("nvdiffrast")
    _stub("nvdiffrast.torch", texture=fn)
    sys.modules["nvdiffrast"].torch = sys.modules["nvdiffrast.torch"]


def load():
    """-> namespace with gaussian_model, sh_utils, general_utils, graphics_utils, nets_lbs, nets_pose (reference
    modules).  render / baking are imported lazily by load_renderer() (they need diff_gaussian_rasterization)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise FileNotFoundError(f"{REF_PY}: reference python modules not installed (oracle/install_ref_py.sh)")
    _stub("knn_cuda", KNN=BruteForceKNN)
    for name in ("plyfile", "trimesh", "imageio", "imageio.v2", "gs_ir", "gs_ir._C"):
        _stub(name)
    sys.modules["plyfile"].PlyData = None
    sys.modules["plyfile"].PlyElement = None
    sys.modules["imageio"].v2 = sys.modules["imageio.v2"]
    sys.modules["gs_ir"]._C = sys.modules["gs_ir._C"]
    if "nvdiffrast.torch" not in sys.modules:
        set_nvdiffrast_texture(None)
    if "simple_knn._C" not in sys.modules:
        _stub("simple_knn")
        _stub("simple_knn._C", distCUDA2=None)
    # our own top-level names that collide with the reference's package names must not shadow them
    for clash in ("utils", "scene", "nets", "smpl", "arguments", "transform", "baking", "gaussian_renderer"):
        m = sys.modules.get(clash)
        if m is not None and not str(getattr(m, "__file__", "") or "").startswith(REF_PY):
            raise ImportError(f"module {clash!r} is already imported from {getattr(m, '__file__', None)}")
    if REF_PY not in sys.path:
        sys.path.insert(0, REF_PY)
    ns = types.SimpleNamespace()
    ns.general_utils = importlib.import_module("utils.general_utils")
    ns.sh_utils = importlib.import_module("utils.sh_utils")
    ns.graphics_utils = importlib.import_module("utils.graphics_utils")
    ns.gaussian_model = importlib.import_module("scene.gaussian_model")
    sys.modules["scene"].Scene = None
    sys.modules["scene"].GaussianModel = ns.gaussian_model.GaussianModel
    ns.nets_lbs = importlib.import_module("nets.mlp_delta_weight_lbs")
    ns.nets_pose = importlib.import_module("nets.mlp_delta_body_pose")
    ns.GaussianModel = ns.gaussian_model.GaussianModel
    _loaded = ns
    return ns


def load_renderer():
    """-> (gaussian_renderer module, baking module) of the reference.  `diff_gaussian_rasterization` must already be in
    sys.modules (the implementation under test)."""
    ns = load()
    if "diff_gaussian_rasterization" not in sys.modules:
        raise ImportError("install a diff_gaussian_rasterization implementation in sys.modules first")
    # gaussian_renderer first: baking imports GaussianModel back from it (the reference's own import order, train.py)
    ns.gaussian_renderer = importlib.import_module("gaussian_renderer")
    ns.baking = importlib.import_module("baking")
    return ns.gaussian_renderer, ns.baking


def unload_renderer():
    """Forget gaussian_renderer / baking (they bind diff_gaussian_rasterization at import): call before switching the
    rasteriser implementation."""
    for name in ("gaussian_renderer", "baking"):
        sys.modules.pop(name, None)


def make_model(smpl, leaves, device, sh_degree=3, motion_offset=False):
    """A reference GaussianModel (scene/gaussian_model.py:32) around synthetic tensors, built without its __init__ (which
    reads SMPL pickles from assets/): fields set exactly as __init__ / create_from_pcd leave them."""
    import torch
    ns = load()
    m = ns.GaussianModel.__new__(ns.GaussianModel)
    m.active_sh_degree = sh_degree
    m.max_sh_degree = sh_degree
    m.setup_functions()
    m.device = torch.device(device)
    m.SMPL_NEUTRAL = {k: v.to(device) for k, v in smpl.items()}
    m.knn = BruteForceKNN(1)
    m.knn_near_2 = BruteForceKNN(2)
    m.knn_near_3 = BruteForceKNN(3)
    m.motion_offset_flag = motion_offset
    m._xyz = leaves["xyz"]
    m._features_dc = leaves["features"][:, :1]
    m._features_rest = leaves["features"][:, 1:]
    m._scaling = leaves["log_scale"]
    m._rotation = leaves["rotation"]
    m._opacity = leaves["opacity_logit"]
    m._normal = leaves["normals"]
    m._albedo = leaves.get("albedo", torch.zeros_like(leaves["xyz"]))
    m._roughness = leaves.get("roughness", torch.zeros_like(leaves["opacity_logit"]))
    m.optimizer = None
    if motion_offset:  # __init__ :94-103: the pose-correction and LBS-weight-offset networks, default initialisation
        torch.manual_seed(0)
        m.pose_decoder = ns.nets_pose.BodyPoseRefiner(total_bones=24, embedding_size=69, mlp_width=128, mlp_depth=2).to(device)
        m.lweight_offset_decoder = ns.nets_lbs.LBSOffsetDecoder(total_bones=24).to(device)
    return m
