#!/bin/bash
# b200jpeg on the 256 files of the C2 workload (host -> host of the C++ surface against bench.py's e2e)
D=/tmp/c2files; mkdir -p $D
python - <<'PY'
import sys; sys.path.insert(0, '.')
import bench
class A: workload, images, quality = "C2", 256, 0
wl = bench.workload(A)
for i, f in enumerate(bench.make_files(wl, 0, 256)):
    open('/tmp/c2files/f%03d.jpg' % i, 'wb').write(f.tobytes())
PY
jpeg_decoder_b200/b200jpeg $D/*.jpg --no-write --repeat 4
