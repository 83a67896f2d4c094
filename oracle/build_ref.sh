#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Builds the UNMODIFIED reference blend library
# (external/multiblend) from the sources where they lie under /root/reference
# into oracle/_ref/libmb_ref.so. Nothing is copied into the repo; outputs go
# only into oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).
# Configuration = xpano's: no MULTIBLEND_WITH_{JPEG,PNG,TIFF}, -msse4.1
# (external/multiblend/CMakeLists.txt:16-18, 89-91), optimisation as a CMake
# Release build has it (-O3 -DNDEBUG).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${XPANO_REFERENCE:-/root/reference}"
MB="$REF/external/multiblend"
OUT="$HERE/_ref"
if [ ! -d "$MB/src" ]; then
  echo "build_ref: $MB not present; keeping prebuilt $OUT (if any)" >&2
  exit 0
fi
SPDLOG=""
for cand in \
  /usr/include \
  /opt/prime-rl/.venv/lib/python3.12/site-packages/flashinfer/data/spdlog/include \
  $(python3 - <<'PY' 2>/dev/null || true
import importlib.util, os
s = importlib.util.find_spec("flashinfer")
if s and s.submodule_search_locations:
    print(os.path.join(list(s.submodule_search_locations)[0], "data/spdlog/include"))
PY
); do
  if [ -f "$cand/spdlog/spdlog.h" ]; then SPDLOG="$cand"; break; fi
done
if [ -z "$SPDLOG" ]; then
  echo "build_ref: spdlog headers not found" >&2
  exit 1
fi
mkdir -p "$OUT"
SRCS=(aligned_ptr flex functions image logging mapalloc multiblend pnger pyramid)
OBJS=()
for s in "${SRCS[@]}"; do
  o="$OUT/$s.o"
  if [ ! -f "$o" ] || [ "$MB/src/$s.cpp" -nt "$o" ]; then
    g++ -std=c++20 -O3 -DNDEBUG -msse4.1 -fPIC -w -c "$MB/src/$s.cpp" -o "$o" \
      -I"$MB/include" -I"$REF/external/simde" \
      -I"$REF/external/thread-pool/include" -I"$SPDLOG" &
  fi
  OBJS+=("$o")
done
wait
g++ -std=c++20 -O3 -DNDEBUG -msse4.1 -fPIC -w -shared "$HERE/ref_driver.cpp" "${OBJS[@]}" \
  -I"$MB/include" -I"$REF/external/simde" \
  -I"$REF/external/thread-pool/include" -I"$SPDLOG" \
  -lpthread -o "$OUT/libmb_ref.so"
echo "build_ref: built $OUT/libmb_ref.so"
