#!/bin/bash
# A/B of alternative core builds: tools/ab_libs.sh "lib1 lib2 ..." "scene:spp ..."
for sc in $2; do
  s=${sc%%:*}; n=${sc##*:}
  for lib in $1; do
    echo -n "$lib  "; RTB_CORE_LIB=$lib python tools/knob_sweep.py $s $n ''
  done
done
