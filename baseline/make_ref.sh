#!/bin/bash
# TEST / BENCH INFRASTRUCTURE -- builds the git-ignored baseline/_ref/ tree that lets `bench.py --impl reference` time
# the UNMODIFIED reference (swami1995/SimuRLacra, Pyrado) on the GPU box's host cores, where /root/reference does not
# exist.  Run in the build container (where /root/reference is mounted):
#
#     bash baseline/make_ref.sh
#
# What it does
#   1. copies the Python sources of /root/reference/Pyrado/pyrado (no meshes / assets: the pysim path needs none) to
#      baseline/_ref/pyrado            -- git-ignored, NOT gpurun-ignored: it travels to the box like a built .so
#   2. copies oracle/ref_shim.py (stubs for the reference's absent third-party imports) next to it and writes a
#      sitecustomize.py that installs the shim in every interpreter started with PYTHONPATH=baseline/_ref and
#      SRE_REF_SHIM=1 (the reference's SamplerPool workers are *spawned*, so they must find the shim on their own)
#   3. applies the two-line patch SURVEY.md 8(c) describes to the COPY of sampling/rollout.py: the fork's rollout()
#      reads `env._wrapped_env._wrapped_env._th_ddot` (lines 238, 311), which only exists for
#      ObsActCat(ActNorm(QCartPole)); every other env chain (QQube + DR of config C2) raises AttributeError.
#      The patch reads the attribute from the inner env if it has one.  Nothing else is touched.
# pip install --target baseline/_ref /root/reference is NOT used: the reference's setup.py pulls dependencies that are
# not in the offline wheelhouse (init_args_serializer, colorama, ...) and builds nothing this path needs.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC=${1:-/root/reference/Pyrado/pyrado}
DST="$HERE/_ref"
[ -d "$SRC" ] || { echo "reference not found at $SRC"; exit 1; }
rm -rf "$DST"
mkdir -p "$DST"
(cd "$(dirname "$SRC")" && find pyrado -name '*.py' -print0 | tar --null -T - -cf - | tar -xf - -C "$DST")
sed -e 's#^REF_ROOT = .*#REF_ROOT = __import__("os").path.dirname(__import__("os").path.abspath(__file__))#' \
    "$HERE/../oracle/ref_shim.py" > "$DST/ref_shim.py"
cat > "$DST/sitecustomize.py" <<'PY'
import os
if os.environ.get("SRE_REF_SHIM") == "1":
    import ref_shim
    ref_shim.install()
PY
R="$DST/pyrado/sampling/rollout.py"
n=$(grep -c 'env\._wrapped_env\._wrapped_env\._th_ddot' "$R")
sed -i 's#env\._wrapped_env\._wrapped_env\._th_ddot#getattr(inner_env(env), "_th_ddot", None)#g' "$R"
echo "baseline/_ref: $(find "$DST" -name '*.py' | wc -l) python files, $(du -sh "$DST" | cut -f1); patched $n occurrences in sampling/rollout.py"
