#!/usr/bin/env bash
# Builds pigs_b200/_Cext.so: the compiled torch binding (torch_ext.cpp) over libpigs_b200.so.  Host compiler only — the
# binding holds no kernel; it links the library by relative rpath so the pair travels together.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../_Cext.so"
PY=${PYTHON:-python}
if [ -f "$OUT" ] && [ "$OUT" -nt "$HERE/torch_ext.cpp" ] && [ "$OUT" -nt "$HERE/../../include/pigs_b200.h" ]; then exit 0; fi
TORCH_DIR=$($PY -c "import torch,os;print(os.path.dirname(torch.__file__))")
PYINC=$($PY -c "import sysconfig;print(sysconfig.get_paths()['include'])")
CUDA=${CUDA_HOME:-/usr/local/cuda}
g++ -O2 -std=c++17 -fPIC -shared -w -DTORCH_EXTENSION_NAME=_Cext -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1 \
    -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include" -I"$PYINC" -I"$CUDA/include" \
    "$HERE/torch_ext.cpp" -o "$OUT" \
    -L"$HERE/.." -l:libpigs_b200.so -L"$TORCH_DIR/lib" -lc10 -lc10_cuda -ltorch_cpu -ltorch_cuda -ltorch -ltorch_python \
    -L"$CUDA/lib64" -lcudart \
    -Wl,-rpath,'$ORIGIN' -Wl,-rpath,"$TORCH_DIR/lib" -Wl,-rpath,"$CUDA/lib64"
echo "[build_ext] built $OUT"
