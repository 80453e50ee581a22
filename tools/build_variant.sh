#!/bin/bash
# usage: tools/build_variant.sh NAME "-DGSR_X=1 ..."   -> build/variants/NAME.so  (A/B builds of the same ABI; select with GSRAST_B200_LIB)
set -e
NAME=$1; FLAGS=${2:-}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p $ROOT/build/variants
make -s -C $ROOT/mvs_gaussian_splatting_b200/csrc -j OBJDIR=$ROOT/build/variants/obj_$NAME OUT=$ROOT/build/variants/$NAME.so EXTRA_NVFLAGS="$FLAGS" > /dev/null
grep -h -A2 "radix_pass_kernelI[jt]Lb[01]ELi[78]" $ROOT/build/variants/obj_$NAME/gsr_binning.ptxas.log | grep -E "registers|spill" | sort | uniq -c | head -8
echo "built build/variants/$NAME.so"
