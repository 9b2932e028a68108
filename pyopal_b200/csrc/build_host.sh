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
