#!/bin/bash
# Build a second copy of the library from another revision of csrc/, for same-board A/B timing with
# profiles/lm_ab.py --lib / profiles/cos_ab.py --lib:    profiles/build_variant.sh <git-rev> <name>
# Output: profiles/ab/<name>/libflr_b200.so (git-ignored; travels to the GPU box).
set -e
cd "$(dirname "$0")/.."
rev=$1; name=$2
pkg=fairness-llm-replication_b200
out=profiles/ab/$name
rm -rf "$out"; mkdir -p "$out/p/csrc" "$out/include"      # csrc includes ../../include/flr_b200.h
for f in $(git ls-tree --name-only "$rev" $pkg/csrc/ include/); do
    case $f in include/*) git show "$rev:$f" > "$out/include/$(basename $f)";; *) git show "$rev:$f" > "$out/p/csrc/$(basename $f)";; esac
done
pids=()
for src in "$out"/p/csrc/*.cu; do
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fopenmp,-ffp-contract=off,-O3 \
         --expt-relaxed-constexpr -I "$out/p/csrc" -I "$out/include" -c "$src" -o "${src%.cu}.o" 2>"${src%.cu}.log" &
    pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -o "$out/libflr_b200.so" "$out"/p/csrc/*.o -Xcompiler -fopenmp -lgomp
echo "$out/libflr_b200.so"
