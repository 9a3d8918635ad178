#!/bin/bash
# The workload meant for compute-sanitizer (tools/sanitize_frames.py: smoke test, every algorithm whole and tiled with the
# light prefetch, the exact variants, a 3-band group run with both exchange modes) on the VR_CHECKED build, in which the
# kernels bounds-check every index of the segment path themselves, plus the GPU parity suite on that build.
# compute-sanitizer itself is closed on this pool; nothing in this repository starts it.
# Build first, in the container: python -c "import __graft_entry__ as g; g.build_checked()"
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
[ -f volumerenderer_b200/libvr_checked.so ] || { echo "libvr_checked.so missing"; exit 2; }
cp volumerenderer_b200/libvr.so /tmp/libvr_product.so
cp volumerenderer_b200/libvr_checked.so volumerenderer_b200/libvr.so     # (the GPU box works on a scratch copy of the repo)
python - > $OUT/${TAG}_checked_selftest.log 2>&1 <<P
import sys; sys.path.insert(0, ".")
import volumerenderer_b200 as vr
from volumerenderer_b200.renderer import VrError
L = vr.load()
print(L.vr_version().decode())
assert b"VR_CHECKED" in L.vr_version()
r = vr.VolumeRenderer(0)
L.vr_checked_selftest(r._ctx)
try:
    r.sync()
    print("selftest FAILED: the violated check was not reported")
    sys.exit(1)
except VrError as e:
    print("selftest ok: a violated device check is reported by vr_sync ->", e)
r.sync()   # the mask is cleared once reported
print("selftest ok: clean afterwards")
P
echo "checked selftest rc=$?"; cat $OUT/${TAG}_checked_selftest.log
python tools/sanitize_frames.py > $OUT/${TAG}_checked_frames.log 2>&1; echo "checked frames rc=$?"; tail -3 $OUT/${TAG}_checked_frames.log
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_checked_pytest.log 2>&1; echo "checked pytest rc=$?"; tail -2 $OUT/${TAG}_checked_pytest.log
cp /tmp/libvr_product.so volumerenderer_b200/libvr.so
