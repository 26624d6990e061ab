#!/usr/bin/env bash
# Builds the UNMODIFIED reference CUDA extension (src/droid.cpp + 3 .cu files) for sm_100a as
# module `droid_backends_ref` into baseline/_ref/ (git-ignored; travels to the GPU box).
# It is used ONLY (a) to generate golden vectors on a B200 (tests/golden/make_ref_cuda_golden.py)
# and (b) as an extra timed baseline in bench.py. Never imported by the product path.
#
# The only edit is applied to a scratch COPY under /tmp: `.type()` -> `.scalar_type()` in the three
# AT_DISPATCH call sites (correlation_kernels.cu:145,176; altcorr_kernel.cu:309), which torch>=2.x
# requires.  No reference source is copied into the repo.
set -euo pipefail
REF=${REF:-/root/reference}
OUT="$(cd "$(dirname "$0")" && pwd)/_ref"
TMP=$(mktemp -d /tmp/refbuild.XXXX)
mkdir -p "$OUT"
cp "$REF"/src/droid.cpp "$REF"/src/droid_kernels.cu "$REF"/src/correlation_kernels.cu "$REF"/src/altcorr_kernel.cu "$TMP"/
sed -i 's/volume\.type()/volume.scalar_type()/; s/fmap1\.type()/fmap1.scalar_type()/' "$TMP"/correlation_kernels.cu "$TMP"/altcorr_kernel.cu
PY=${PY:-python}
read -r TINC PYINC TLIB EXT <<<"$($PY - <<'PYEOF'
import sysconfig, torch.utils.cpp_extension as c, os
print(" ".join("-I"+p for p in c.include_paths("cuda")).replace(" ", ","), "-I"+sysconfig.get_paths()["include"], os.path.join(os.path.dirname(c.__file__), "..", "lib"), sysconfig.get_config_var("EXT_SUFFIX"))
PYEOF
)"
TINC=${TINC//,/ }
COMMON="-DTORCH_EXTENSION_NAME=droid_backends_ref -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1 $TINC $PYINC -I$REF/thirdparty/eigen -std=c++17 -O3"
NV="nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo --expt-relaxed-constexpr -Xcompiler -fPIC $COMMON"
cd "$TMP"
for f in droid_kernels correlation_kernels altcorr_kernel; do
  $NV -c $f.cu -o $f.o &
done
g++ -fPIC $COMMON -c droid.cpp -o droid.o &
wait
g++ -shared -o "$OUT/droid_backends_ref$EXT" droid.o droid_kernels.o correlation_kernels.o altcorr_kernel.o \
  -L"$TLIB" -ltorch -ltorch_cpu -ltorch_cuda -ltorch_python -lc10 -lc10_cuda -L/usr/local/cuda/lib64 -lcudart
ls -la "$OUT"
rm -rf "$TMP"
