#!/bin/sh
# Syntax-only compile of the PD_BUILD_OP shim against the reference's own headers
# (paddle/extension.h and its closure), with stand-ins under stubs/ for the two
# build-generated headers (api.h, sparse_api.h), the generated-API re-export
# (tensor_compat.h) and the un-vendored gflags.  Paddle itself is not installed here,
# so this is as far as the shim can be checked on the authoring box.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${CTRFE_REFERENCE_ROOT:-/root/reference}"
g++ -std=c++17 -fsyntax-only -w -DPADDLE_WITH_CUDA -I "$HERE/stubs" -I "$REF" -I "$HERE/../../include" \
    -I /usr/local/cuda/include "$HERE/ctrfe_paddle_ops.cc"
echo "ctrfe_paddle_ops.cc: syntax OK against $REF"
