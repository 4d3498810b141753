#!/bin/bash
# A/B of conv-kernel build variants: each variant's library is built beforehand into cmf_b200/lib/variants/<name>.so
for v in "$@"; do
  cp cmf_b200/lib/variants/$v.so cmf_b200/lib/libcmf_b200.so
  echo "== $v"
  timeout 120 python scripts/one_conv_bench.py
done
