#!/bin/sh
# Builds tests/seam/_build/ref_seam: the REFERENCE's own parser feeding libb200jpeg.so through fill_desc()
# (see ref_seam_post.h).  Needs the reference checkout (REF, default /root/reference): built in the build container,
# shipped prebuilt to the GPU box like oracle/_ref/ (git-ignored, not gpurun-ignored).  No reference source is copied.
set -e
cd "$(dirname "$0")/../.."
REF=${REF:-/root/reference}
if [ ! -f "$REF/jpeg.cpp" ]; then echo "reference sources absent at $REF: keeping prebuilt tests/seam/_build/ref_seam"; exit 0; fi
mkdir -p tests/seam/_build
( echo '#include "ref_wrapper_pre.h"'; echo '#line 1 "jpeg.cpp"'
  sed -e 's|^#include "include/display.hpp".*|#include "include/types.hpp"\n#include "ref_nodisplay.h"|' "$REF/jpeg.cpp"
  echo; echo '#include "ref_seam_post.h"' ) | \
g++ -O2 -std=c++17 -w -U_FORTIFY_SOURCE -D_FORTIFY_SOURCE=0 -ffp-contract=off -Ioracle -Itests/seam -Iinclude -I"$REF" -x c++ - \
    -o tests/seam/_build/ref_seam -Ljpeg_decoder_b200 -lb200jpeg -Wl,-rpath,'$ORIGIN/../../../jpeg_decoder_b200' -lpthread
echo "built tests/seam/_build/ref_seam"
# CPU-tier variant against the emulation library (tests/emul/_build, built by tests/emul/build.sh)
if [ -f tests/emul/_build/libb200jpeg_emul.so ]; then
( echo '#include "ref_wrapper_pre.h"'; echo '#line 1 "jpeg.cpp"'
  sed -e 's|^#include "include/display.hpp".*|#include "include/types.hpp"\n#include "ref_nodisplay.h"|' "$REF/jpeg.cpp"
  echo; echo '#include "ref_seam_post.h"' ) | \
g++ -DSEAM_EMU -O2 -std=c++17 -w -U_FORTIFY_SOURCE -D_FORTIFY_SOURCE=0 -ffp-contract=off -Ioracle -Itests/seam -Iinclude -I"$REF" -x c++ - \
    -o tests/seam/_build/ref_seam_emu -Ltests/emul/_build -lb200jpeg_emul -Wl,-rpath,'$ORIGIN/../../emul/_build' -lpthread
echo "built tests/seam/_build/ref_seam_emu"
fi
