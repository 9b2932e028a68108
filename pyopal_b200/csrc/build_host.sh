#!/bin/bash
# Build the host-side CPython extension _results (result storage + bulk list constructor) in-tree.
set -e
cd "$(dirname "$0")"
PY=${PYTHON:-python}
EXT=$($PY -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
INC=$($PY -c "import sysconfig; print(sysconfig.get_paths()['include'])")
if [ ! -f ../_results$EXT ] || [ _results.c -nt ../_results$EXT ]; then
  gcc -O2 -fPIC -shared -Wall -I"$INC" _results.c -o ../_results$EXT
fi
echo built $(pwd)/../_results$EXT
# Cython binding of the C-ABI (platform/_cuda_cy.pyx): the reference's host layer is Cython (platform/pyx.in);
# linked against libpyopal_cuda.so next to the package ($ORIGIN/..)
if $PY -c "import Cython" 2>/dev/null; then
  if [ ! -f ../platform/_cuda_cy$EXT ] || [ ../platform/_cuda_cy.pyx -nt ../platform/_cuda_cy$EXT ] || [ ../../include/pyopal_cuda.h -nt ../platform/_cuda_cy$EXT ]; then
    $PY -m cython -3 ../platform/_cuda_cy.pyx -o build/_cuda_cy.c
    gcc -O2 -fPIC -shared -Wno-unused-function -I"$INC" -I../../include build/_cuda_cy.c -o ../platform/_cuda_cy$EXT \
        -L.. -lpyopal_cuda -Wl,-rpath,'$ORIGIN/..'
  fi
  echo built $(pwd)/../platform/_cuda_cy$EXT
fi
