#!/bin/bash
# Copies the reference's own PYTHON modules of the render hot path, unmodified, from where they lie under
# /root/reference into baseline/_ref/ (git-ignored, NOT gpurun-ignored: it travels to the GPU box like oracle/_ref).
# Only there can bench.py's reference arm and tests/test_dropin_reference_gpu.py import the reference's own
# coarse_deform_c2source / get_covariance / eval_sh / render() / bake_set.
# Nothing under baseline/_ref is ever committed; the repo's history holds no reference source.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${GAUHUMAN_REFERENCE:-/root/reference}"
OUT="$HERE/../baseline/_ref"
if [ ! -f "$REF/scene/gaussian_model.py" ]; then
  echo "install_ref_py: $REF not present; keeping any existing baseline/_ref" >&2
  exit 0
fi
mkdir -p "$OUT"/{scene,utils,nets,smpl,gaussian_renderer,arguments,smplx}
cp "$REF/scene/gaussian_model.py" "$OUT/scene/"
cp "$REF"/utils/{general_utils,sh_utils,graphics_utils,system_utils,loss_utils,image_utils}.py "$OUT/utils/"
cp "$REF"/nets/{mlp_delta_body_pose,mlp_delta_weight_lbs,brdf_network}.py "$OUT/nets/"
cp "$REF/smpl/smpl_numpy.py" "$OUT/smpl/"
cp "$REF/gaussian_renderer/__init__.py" "$OUT/gaussian_renderer/"
cp "$REF/arguments/__init__.py" "$OUT/arguments/"
cp "$REF"/smplx/{lbs,utils}.py "$OUT/smplx/"
cp "$REF/transform.py" "$REF/baking.py" "$OUT/"
# package markers (ours, empty): scene/__init__.py of the reference pulls in the dataset readers and is NOT copied
for d in scene utils nets smpl smplx; do : > "$OUT/$d/__init__.py"; done
echo "installed reference python modules into $OUT"
# the reference's own Python front-end of its rasteriser (autograd function, settings tuple, module); its compiled _C
# half is provided on the GPU box by oracle/ref_dgr.py on top of oracle/_ref (the unmodified CUDA sources)
mkdir -p "$OUT/ref_dgr/diff_gaussian_rasterization"
cp "$REF/submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py" "$OUT/ref_dgr/diff_gaussian_rasterization/"
