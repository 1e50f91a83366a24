#!/bin/bash
# tools/variant.sh <name> "<nvcc -D flags>" [tu ...] -- kernel experiment: recompiles the given translation units (default: the
# QQube rollout + the stand-alone tensor-core policy) with extra flags and links them with the objects of the default build into
# simurlacra_b200/lib/libsre_<name>.so (select with SRE_B200_LIB=simurlacra_b200/lib/libsre_<name>.so).
set -eu
cd "$(dirname "$0")/../simurlacra_b200/csrc"
name=$1; flags=$2; shift 2
tus=${*:-"sre_rollout_qq sre_policy_tc"}
make -s -j8 >/dev/null
mkdir -p _build_$name
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr"
pids=""
for t in $tus; do $NV $flags -c $t.cu -o _build_$name/$t.o & pids="$pids $!"; done
for p in $pids; do wait $p; done
objs=""
for o in _build/*.o; do b=$(basename $o); if [ -f _build_$name/$b ]; then objs="$objs _build_$name/$b"; else objs="$objs $o"; fi; done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/libsre_$name.so $objs -cudart static
echo "built simurlacra_b200/lib/libsre_$name.so"
