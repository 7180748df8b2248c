#!/bin/bash
# TEST INFRASTRUCTURE ONLY -- makes the UNMODIFIED reference importable on the GPU box.
# /root/reference does not exist there, and the reference is a flat Python repo without a setup.py (pip cannot install
# it), so this recipe copies exactly the files the manifold_dimension path imports into oracle/_ref/, which is
# git-ignored (no reference source enters the history) but NOT gpurun-ignored (it travels like the built .so).
# Used by: bench.py --impl reference and the cpu_baseline legs (kind "reference"); oracle/ref_shims.py falls back to it
# when /root/reference is absent.  Run by __graft_entry__.build() whenever /root/reference is present.
set -euo pipefail
SRC="${1:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
DST="$HERE/_ref"
FILES="dim_reduction.py sde_lib.py losses.py
lightning_data_modules/__init__.py lightning_data_modules/utils.py lightning_data_modules/KSphereDataset.py
lightning_modules/__init__.py lightning_modules/utils.py lightning_modules/BaseSdeGenerativeModel.py
lightning_modules/KSphereGroundTruthModel.py
models/__init__.py models/utils.py models/fcn.py models/ksphere_gt.py models/ddpm.py models/layers.py
models/normalization.py models/ncsnpp.py models/layerspp.py models/up_or_down_sampling.py models/BeatGANsUNET.py
models/BeatGANsblocks.py models/BeatGANs_nn.py models/BeatGANs_choices.py models/config_base.py
op/__init__.py op/fused_act.py op/upfirdn2d.py
sampling/__init__.py sampling/correctors.py sampling/predictors.py sampling/unconditional.py"
[ -d "$SRC" ] || { echo "build_ref: $SRC not found (nothing to do on the GPU box)"; exit 0; }
rm -rf "$DST"
for f in $FILES; do
  mkdir -p "$DST/$(dirname "$f")"
  cp "$SRC/$f" "$DST/$f"
done
( cd "$SRC" && sha256sum $FILES ) > "$DST/SHA256SUMS"
echo "build_ref: $(echo $FILES | wc -w) files -> $DST"
