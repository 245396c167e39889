#!/bin/bash
# Developer tool: builds A/B variants of the product library into hpg-fastq_b200/csrc/build_<name>/libhpgq_cuda.so
# (select one at run time with HPGQ_LIB=<path>).  usage: tools/build_variants.sh name1="-DFLAG=0" name2="-DX=1 -DY=0"
cd "$(dirname "$0")/../hpg-fastq_b200/csrc" || exit 1
for spec in "$@"; do
  name="${spec%%=*}"; flags="${spec#*=}"
  echo "== variant $name : $flags"
  make -j8 BUILD="build_$name" OUT="$PWD/build_$name/libhpgq_cuda.so" VARIANT="$flags" "$PWD/build_$name/libhpgq_cuda.so" 2>&1 | grep -E "error|Li5ELi8ELb0|spill stores" | grep -v " 0 bytes spill" | head
done
