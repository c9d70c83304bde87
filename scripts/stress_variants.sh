#!/bin/bash
# scripts/stress_variants.sh N tag1 tag2 ... : scripts/stress.sh with each variant library swapped in
N=$1; shift
cp p-ntfields_b200/pntf/libpntf_b200.so /tmp/libpntf_orig.so
for T in "$@"; do
  cp p-ntfields_b200/pntf/variants/libpntf_b200.$T.so p-ntfields_b200/pntf/libpntf_b200.so
  echo "== $T"; bash scripts/stress.sh $N 2>/dev/null | tail -3
done
cp /tmp/libpntf_orig.so p-ntfields_b200/pntf/libpntf_b200.so
