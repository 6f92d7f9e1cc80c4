#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the UNMODIFIED reference CUDA extension
# (kr4b/pigs, diff-gaussian-sampling) from the sources where they lie under
# /root/reference into oracle/_ref/_C.so.  Nothing is copied into the repo;
# oracle/_ref/ is git-ignored but travels to the GPU box with the snapshot.
#
# The reference's own build system (setup.py / CUDAExtension) is NOT run: this
# is a direct nvcc recipe over its six source files.  The resulting pybind
# module `_C` exports the reference's 12 entry points (ext.cpp:20-31) and is
# used ONLY by tests/ and bench.py as the bit-level oracle / GPU baseline.
set -euo pipefail
REF=${REF:-/root/reference/diff-gaussian-sampling}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "[build_ref] $REF not present (GPU box?) — using prebuilt $OUT/_C.so if any"
  exit 0
fi
mkdir -p "$OUT/obj"
PY=${PYTHON:-python}
TORCH_DIR=$($PY -c "import torch,os;print(os.path.dirname(torch.__file__))" 2>/dev/null)
PYINC=$($PY -c "import sysconfig;print(sysconfig.get_paths()['include'])")
INC="-I$TORCH_DIR/include -I$TORCH_DIR/include/torch/csrc/api/include -I$PYINC -I$REF/third_party/glm -I$REF"
DEFS="-DTORCH_EXTENSION_NAME=_C -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1"
# same optimisation level the reference's setup.py yields (torch default -O2... nvcc default -O3 device)
NVFLAGS="-gencode arch=compute_100a,code=sm_100a -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC -Xcompiler -fno-gnu-unique -w $DEFS $INC"
SRCS="cuda_sampler/sampler_impl.cu cuda_sampler/forward.cu cuda_sampler/backward.cu aggregate_neighbors.cu sample_points.cu ext.cpp"
pids=()
objs=()
for s in $SRCS; do
  o="$OUT/obj/$(basename ${s%.*}).o"
  objs+=("$o")
  if [ ! -f "$o" ] || [ "$REF/$s" -nt "$o" ]; then
    ( nvcc $NVFLAGS -x cu -c "$REF/$s" -o "$o" ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
nvcc -shared -o "$OUT/_C.so" "${objs[@]}" -L"$TORCH_DIR/lib" -lc10 -lc10_cuda -ltorch_cpu -ltorch_cuda -ltorch -ltorch_python \
  -Xlinker -rpath -Xlinker "$TORCH_DIR/lib"
echo "[build_ref] built $OUT/_C.so"
