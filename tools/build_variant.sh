#!/bin/bash
# usage: tools/build_variant.sh NAME [-DMACRO=VALUE ...]  ->  nusiprop_b200/lib/variants/libnusiprop_b200_NAME.so
# (tuning experiments: the same sources with other compile-time parameters; timed by tools/gpu_variant_time.py)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
python - "$name" "$@" <<'P'
import os, sys
sys.path.insert(0, ".")
from nusiprop_b200.build import build_library, _HERE
out = os.path.join(_HERE, "lib", "variants", "libnusiprop_b200_%s.so" % sys.argv[1])
os.makedirs(os.path.dirname(out), exist_ok=True)
print(build_library(extra_flags=sys.argv[2:], out=out))
P
