#!/bin/bash
# A/B of TRACK loop parameters on the C3 frame: each variant is a full build of csrc/ with one -D (build/ab/<variant>), selected with LR_B200_LIB.
for v in base r8 r24 s10; do
  if [ $v = base ]; then unset LR_B200_LIB; else export LR_B200_LIB=$PWD/build/ab/$v/liverrenderer_b200/csrc/liblr_b200.so; fi
  for i in 1 2; do echo -n "$v: "; python tools/prof_render.py c3 1080 256 dda 3 2>&1 | tail -1; done
done
