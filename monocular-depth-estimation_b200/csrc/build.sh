#!/usr/bin/env bash
# Build libswin_attn.so for sm_100a, in-tree (the .so travels with the repo to the GPU box).
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
root="$(cd "$here/../.." && pwd)"
out="$here/../libswin_attn.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
flags=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC
       -I"$root/include" -I"$here" -Xptxas -v)
# SWIN_TC_TRACE=1: compile the per-pair clock stamps of the tcgen05 kernels in (tools/tc_trace.py); off in the product build
[ "${SWIN_TC_TRACE:-0}" = "1" ] && flags+=(-DSWIN_TC_TRACE)
mkdir -p "$here/build"
pids=()
for f in swin_attn_simt swin_attn_tc swin_attn_tc12 swin_attn_mma12 swin_attn_tc_bwd swin_linear_tc swin_block_ops mha_global mha_tc swin_attn_api; do
  "$NVCC" "${flags[@]}" -c "$here/$f.cu" -o "$here/build/$f.o" > "$here/build/$f.log" 2>&1 &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p" || { cat "$here"/build/*.log; exit 1; }; done
"$NVCC" -shared -o "$out" "$here"/build/swin_attn_simt.o "$here"/build/swin_attn_tc.o "$here"/build/swin_attn_tc12.o "$here"/build/swin_attn_mma12.o "$here"/build/swin_attn_tc_bwd.o "$here"/build/swin_linear_tc.o "$here"/build/swin_block_ops.o "$here"/build/mha_global.o "$here"/build/mha_tc.o "$here"/build/swin_attn_api.o -lcudart
echo "built $out"

# torch.library binding (C++ dispatcher ops over the C ABI): host-only code, g++, links libswin_attn.so by $ORIGIN
PY="${PYTHON:-python}"
tout="$here/../libswin_attn_torch.so"
if [ "$here/torch_binding.cpp" -nt "$tout" ] || [ "$root/include/swin_attn.h" -nt "$tout" ] || [ ! -f "$tout" ]; then
  read -r tinc tlib tabi < <("$PY" - <<'PYEOF'
import os, torch
d = os.path.dirname(torch.__file__)
print(os.path.join(d, "include"), os.path.join(d, "lib"), int(torch._C._GLIBCXX_USE_CXX11_ABI))
PYEOF
)
  g++ -O2 -std=c++17 -fPIC -shared -D_GLIBCXX_USE_CXX11_ABI="$tabi" -DUSE_CUDA \
      -I"$tinc" -I"$tinc/torch/csrc/api/include" -I/usr/local/cuda/include \
      "$here/torch_binding.cpp" -o "$tout" \
      -L"$here/.." -lswin_attn -L"$tlib" -lc10 -lc10_cuda -ltorch_cpu -ltorch_cuda -ltorch -L/usr/local/cuda/lib64 -lcudart \
      -Wl,-rpath,'$ORIGIN' -Wl,-rpath,"$tlib" > "$here/build/torch_binding.log" 2>&1 || { cat "$here/build/torch_binding.log"; exit 1; }
fi
echo "built $tout"
