#!/bin/bash
# Build libslate_b200.so: the host C++ mirror of SLATE's interface, above the C ABI.
set -e
cd "$(dirname "$0")/.."
H=slate_d35_b200/host
/usr/bin/g++ -std=c++17 -O2 -fPIC -Wall -Iinclude -I$H/include -shared \
    -o slate_d35_b200/lib/libslate_b200.so $H/src/*.cc \
    -Lslate_d35_b200/lib -ltile_engine -Wl,-rpath,'$ORIGIN'
# Fortran-name interposition library (pdgemm_, pdpotrf_, ...) over libslate_b200.so
/usr/bin/g++ -std=c++17 -O2 -fPIC -Wall -Iinclude -shared \
    -o slate_d35_b200/lib/libslate_b200_scalapack.so $H/scalapack_names/scalapack_names.cc \
    -Lslate_d35_b200/lib -lslate_b200 -Wl,-rpath,'$ORIGIN'
