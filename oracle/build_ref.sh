#!/usr/bin/env bash
# TEST/BASELINE INFRASTRUCTURE — not part of the product.
#
# Compiles the UNMODIFIED reference (leejaeyong7/tiny-cuda-nn, /root/reference) CUDA sources
# where they lie, directly with nvcc (the reference's own CMake build is NOT run), for sm_100,
# and links them with oracle/ref_driver.cu (our own small harness over the reference's public
# C++ API) into oracle/_ref/ref_driver.  Outputs go ONLY to oracle/_ref/ (git-ignored; travels
# to the GPU box with gpurun).  The reference has no CPU path, so the binary can only run on
# the GPU box; there it is (1) the correctness oracle that produces golden dumps and (2) the
# "reference CUDA build on one B200" baseline of BASELINE.md R1.
#
# Source list = CMakeLists.txt:249-263 of the reference; flags = CMakeLists.txt:205-222.
# The arch clamp of CMakeLists.txt:143-166 (max 90) is bypassed by passing the gencode directly
# (SURVEY.md F1, §8c route (2)).
#
# usage: oracle/build_ref.sh [jobs]        (background it: fully_fused_mlp.cu takes very long)
set -u
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
JOBS=${1:-4}
mkdir -p "$OUT/obj"
if [ ! -d "$REF/src" ]; then echo "no reference at $REF — using prebuilt $OUT if any"; exit 0; fi

NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-std=c++14 -O3 -gencode arch=compute_100,code=sm_100 -DTCNN_MIN_GPU_ARCH=100 -DFMT_HEADER_ONLY
       --extended-lambda --expt-relaxed-constexpr --split-compile ${SPLIT:-3}
       -Xcompiler=-Wno-float-conversion -Xcompiler=-fno-strict-aliasing -Xcompiler=-fPIC
       -Xcudafe=--diag_suppress=unrecognized_gcc_pragma -w
       -I"$REF/include" -I"$REF/dependencies" -I"$REF/dependencies/cutlass/include"
       -I"$REF/dependencies/cutlass/tools/util/include" -I"$REF/dependencies/fmt/include")

SRCS=(common_host cpp_api cutlass_mlp encoding loss network object optimizer reduce_sum fully_fused_mlp)

compile_one() {
  local s=$1
  local o="$OUT/obj/$s.o"
  if [ -f "$o" ] && [ "$o" -nt "$REF/src/$s.cu" ]; then echo "[ref] $s.o up to date"; return 0; fi
  local t0=$(date +%s)
  "$NVCC" "${FLAGS[@]}" -c "$REF/src/$s.cu" -o "$o.tmp" > "$OUT/obj/$s.log" 2>&1 && mv "$o.tmp" "$o"
  local rc=$?
  echo "[ref] $s.cu rc=$rc $(( $(date +%s) - t0 ))s"
  return $rc
}

# fully_fused_mlp first (longest), then the rest, JOBS at a time
pids=()
running=0
for s in fully_fused_mlp "${SRCS[@]:0:9}"; do
  compile_one "$s" &
  pids+=($!)
  running=$((running+1))
  if [ "$running" -ge "$JOBS" ]; then wait -n; running=$((running-1)); fi
done
wait
for s in "${SRCS[@]}"; do [ -f "$OUT/obj/$s.o" ] || { echo "[ref] missing $s.o"; exit 1; }; done

# driver (ours) + link
"$NVCC" "${FLAGS[@]}" -c "$HERE/ref_driver.cu" -o "$OUT/obj/ref_driver.o" > "$OUT/obj/ref_driver.log" 2>&1 || { echo "[ref] driver compile failed"; tail -30 "$OUT/obj/ref_driver.log"; exit 1; }
objs=(); for s in "${SRCS[@]}"; do objs+=("$OUT/obj/$s.o"); done
"$NVCC" -gencode arch=compute_100,code=sm_100 "${objs[@]}" "$OUT/obj/ref_driver.o" -o "$OUT/ref_driver" -lcuda -Xlinker -rpath=/usr/local/cuda/lib64 || { echo "[ref] link failed"; exit 1; }
echo "[ref] built $OUT/ref_driver"
