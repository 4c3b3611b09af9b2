#!/bin/bash
# A/B on one box: bench4 with the built libraries, then with each variants/<suffix>/lib*.so swapped in (scripts/build_variant.sh)
# usage: ab_bench.sh <suffix> [<suffix> ...]
S=robot_mpcs_b200/solvers
mkdir -p variants/_built
for d in $S/*/; do m=$(basename $d); cp $d/lib$m.so variants/_built/; done
put() { for f in variants/$1/lib*.so; do m=$(basename $f .so); m=${m#lib}; cp $f $S/$m/lib$m.so; done; }
echo "== built"; bash scripts/bench4.sh
for v in "$@"; do put $v; echo "== $v"; bash scripts/bench4.sh; done
put _built; echo "== built again"; bash scripts/bench4.sh
