#!/usr/bin/env bash
# Build the UNMODIFIED-IN-ARITHMETIC reference (Omegastick/pytorch-cpp-rl) as a CPU oracle.
#
# TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path.
#
# Reads sources where they lie under /root/reference, patches a scratch copy in
# $BUILD (outside the repo) for the 5 libtorch-1.4 -> 2.x API sites listed in
# SURVEY.md §8c (none touches arithmetic), renames `namespace cpprl` to
# `cpprl_ref` with -Dcpprl=cpprl_ref so it can live in one process with the
# product's own `cpprl::` classes, and writes ONLY binaries into oracle/_ref/:
#   oracle/_ref/libcpprl_ref.so   reference classes + oracle/ref_driver.cpp (C ABI for ctypes)
#   oracle/_ref/cpprl_ref_tests   the reference's own doctest suite (19 cases / 297 assertions)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${CPPRL_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
BUILD="${CPPRL_REF_BUILD:-/tmp/cpprl_ref_build}"
PY="${PYTHON:-python}"
if [ ! -d "$REF/src" ]; then echo "reference not present at $REF; keeping prebuilt $OUT" >&2; exit 0; fi
TORCH="$($PY -c 'import torch,os;print(os.path.dirname(torch.__file__))' 2>/dev/null | tail -1)"
mkdir -p "$OUT" "$BUILD"
rm -rf "$BUILD/src" && cp -r "$REF/src" "$BUILD/src"

# --- API-site patches (libtorch 1.4 -> 2.x); arithmetic untouched -------------
sed -i 's/\.index(indices\[index\])/.index({indices[index]})/g' "$BUILD/src/generators/feed_forward_generator.cpp"
sed -i 's/gru_output\.output/std::get<0>(gru_output)/g; s/gru_output\.state/std::get<1>(gru_output)/g' "$BUILD/src/model/nn_base.cpp"
sed -i 's/AT_CHECK/TORCH_CHECK/g' "$BUILD/src/model/model_utils.cpp"
sed -i 's/optimizer->options\.learning_rate(original_learning_rate \* decay_level);/for (auto \&g : optimizer->param_groups()) static_cast<torch::optim::AdamOptions \&>(g.options()).lr(original_learning_rate * decay_level);/' "$BUILD/src/algorithms/ppo.cpp"
sed -i 's/optimizer->options\.learning_rate(original_learning_rate \* decay_level);/for (auto \&g : optimizer->param_groups()) static_cast<torch::optim::RMSpropOptions \&>(g.options()).lr(original_learning_rate * decay_level);/' "$BUILD/src/algorithms/a2c.cpp"
# test-harness-only: randint defaults to Long since libtorch 1.5
sed -i 's/torch::randint(0, 2, {2, 1})/torch::randint(0, 2, {2, 1}, torch::kFloat)/g' "$BUILD/src/algorithms/ppo.cpp" "$BUILD/src/algorithms/a2c.cpp"

CXXFLAGS="-O2 -std=c++17 -fPIC -w -I$REF/include -I$BUILD/src -I$TORCH/include -I$TORCH/include/torch/csrc/api/include -D_GLIBCXX_USE_CXX11_ABI=1"
LDFLAGS="-L$TORCH/lib -Wl,-rpath,$TORCH/lib -ltorch -ltorch_cpu -lc10"
SRCS=$(cd "$BUILD/src" && ls *.cpp algorithms/*.cpp distributions/*.cpp generators/*.cpp model/*.cpp)

mkdir -p "$BUILD/obj_lib" "$BUILD/obj_test"
build_one() {  # $1 = src, $2 = objdir, rest = extra flags
  local src="$1" dir="$2"; shift 2
  local obj="$dir/$(echo "$src" | tr '/' '_').o"
  g++ $CXXFLAGS "$@" -c "$BUILD/src/$src" -o "$obj"
}
export -f build_one; export CXXFLAGS BUILD
JOBS="${JOBS:-$(nproc)}"
# library objects: tests compiled out, namespace renamed
echo "$SRCS" | tr ' ' '\n' | xargs -P "$JOBS" -I{} bash -c 'build_one {} '"$BUILD"'/obj_lib -DDOCTEST_CONFIG_DISABLE -Dcpprl=cpprl_ref'
g++ $CXXFLAGS -Dcpprl=cpprl_ref -I"$HERE" -c "$HERE/ref_driver.cpp" -o "$BUILD/obj_lib/ref_driver.o"
g++ -shared -o "$OUT/libcpprl_ref.so" "$BUILD"/obj_lib/*.o $LDFLAGS
echo "built $OUT/libcpprl_ref.so"

if [ "${CPPRL_REF_TESTS:-1}" = "1" ]; then
  echo "$SRCS" | tr ' ' '\n' | xargs -P "$JOBS" -I{} bash -c 'build_one {} '"$BUILD"'/obj_test'
  g++ $CXXFLAGS -DDOCTEST_CONFIG_NO_POSIX_SIGNALS -c "$BUILD/src/third_party/doctest.cpp" -o "$BUILD/obj_test/doctest_main.o"
  g++ -o "$OUT/cpprl_ref_tests" "$BUILD"/obj_test/*.o $LDFLAGS
  echo "built $OUT/cpprl_ref_tests"
fi
