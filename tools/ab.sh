#!/bin/bash
# usage: ab.sh v1 v2 ... ; runs tools/profile_build.py for each variant lib (and main last)
for v in "$@"; do echo "== $v"; KMI_LIB=$PWD/kmerind_b200/libkmi_$v.so python tools/profile_build.py 2>&1 | tail -1; done
echo "== main"; python tools/profile_build.py 2>&1 | tail -1
