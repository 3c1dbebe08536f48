#!/usr/bin/env bash
# Builds tests/cpp/host_shim_test (the C++ drop-in against the compiled reference).  Needs lib/libcpprl_host.so
# (python -m pytorch_cpp_rl_b200.build --host) and oracle/_ref/libcpprl_ref.so (oracle/build_ref.sh).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
TORCH="$(python -c 'import torch,os;print(os.path.dirname(torch.__file__))' 2>/dev/null | tail -1)"
LIB="$ROOT/pytorch_cpp_rl_b200/lib"
g++ -O1 -std=c++17 -w -D_GLIBCXX_USE_CXX11_ABI=1 -I"$ROOT/include" -I"$ROOT" -I"$TORCH/include" \
    -I"$TORCH/include/torch/csrc/api/include" -I/usr/local/cuda/include "$HERE/host_shim_test.cpp" -o "$HERE/host_shim_test" \
    -L"$LIB" -lcpprl_host -lcpprl_b200 -L"$ROOT/oracle/_ref" -lcpprl_ref -L"$TORCH/lib" -ltorch -ltorch_cpu -lc10 -lc10_cuda \
    -ltorch_cuda -Wl,--no-as-needed -Wl,-rpath,"$LIB" -Wl,-rpath,"$ROOT/oracle/_ref" -Wl,-rpath,"$TORCH/lib"
echo "built $HERE/host_shim_test"
