#!/bin/bash
# Build the csrc/ of a git revision into probtalk_b200/lib_<name>/libprobtalk_b200.so for A/B runs:
#   tools/build_variant.sh <name> [<rev>]        (PROBTALK_B200_LIB=probtalk_b200/lib_<name>/libprobtalk_b200.so selects it)
set -e
name=$1; rev=${2:-HEAD}
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d)
git -C "$root" archive "$rev" probtalk_b200/csrc include | tar -x -C "$tmp"
out="$root/probtalk_b200/lib_$name"; mkdir -p "$out"
objs=""
for f in "$tmp"/probtalk_b200/csrc/*.cu; do
  o="$tmp/$(basename "${f%.cu}").o"
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -c "$f" -o "$o" &
  objs="$objs $o"
done
wait
nvcc -shared -o "$out/libprobtalk_b200.so" $objs -lcudart
rm -rf "$tmp"; echo "$out/libprobtalk_b200.so"
