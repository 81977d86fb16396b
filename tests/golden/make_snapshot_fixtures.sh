#!/bin/bash
# Regenerates tests/golden/ref_nlohmann.msgpack, ref_child.json, ref_parent.json with the REFERENCE's vendored nlohmann
# json (oracle/ref_harness/make_snapshot.cpp).  Dev container only (/root/reference); the fixtures are committed.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${NRF_REFERENCE_DIR:-/root/reference}"
mkdir -p "$HERE/../../oracle/_ref"
g++ -std=c++17 -O1 -w -I"$R/dependencies/tiny-cuda-nn/dependencies" "$HERE/../../oracle/ref_harness/make_snapshot.cpp" \
  -o "$HERE/../../oracle/_ref/make_snapshot"
"$HERE/../../oracle/_ref/make_snapshot" "$HERE/ref_nlohmann.msgpack" "$HERE/ref_child.json" "$HERE/ref_parent.json"
ls -la "$HERE"/ref_nlohmann.msgpack "$HERE"/ref_child.json "$HERE"/ref_parent.json
