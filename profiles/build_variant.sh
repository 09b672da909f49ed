#!/bin/bash
# Build a second copy of the library from another version of csrc/list_metrics.cu, for same-board A/B timing
# with profiles/lm_ab.py --lib:    profiles/build_variant.sh <git-rev | path/to/list_metrics.cu> <name>
# Output: profiles/ab/<name>/libflr_b200.so (git-ignored; travels to the GPU box).
set -e
cd "$(dirname "$0")/.."
src=$1; name=$2
pkg=fairness-llm-replication_b200
out=profiles/ab/$name
mkdir -p "$out"
if [ -f "$src" ]; then cp "$src" "$out/list_metrics.cu"; else git show "$src:$pkg/csrc/list_metrics.cu" > "$out/list_metrics.cu"; fi
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fopenmp,-ffp-contract=off,-O3 \
     --expt-relaxed-constexpr -I $pkg/csrc -I include -c "$out/list_metrics.cu" -o "$out/list_metrics.o"
objs=$(ls $pkg/build/*.o | grep -v list_metrics.o)
nvcc -shared -o "$out/libflr_b200.so" $objs "$out/list_metrics.o" -Xcompiler -fopenmp -lgomp
echo "$out/libflr_b200.so"
