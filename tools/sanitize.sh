#!/bin/bash
# compute-sanitizer memcheck + racecheck over small wavefront renders of the three scene classes (BVH scene, list scene,
# list scene with media), the closest-hit harness and the multi-device entry.  usage: tools/sanitize.sh <tag>
tag=${1:-r02}
cat > /tmp/san_run.py <<'PY'
import sys, numpy as np
sys.path.insert(0, ".")
import __graft_entry__ as e
pkg = e.load_package(); abi = pkg.abi
for name in ("random_ned", "cornell_ned", "cornell_smoke_ned"):
    host = pkg.HostScene(name, seed=12345, wrap=True)
    dev = pkg.DeviceScene(host.desc, device=0)
    W, H, spp = 96, 54, 6
    cam = pkg.main_camera(W, H)
    for kernel in (abi.KERNEL_WAVEFRONT, abi.KERNEL_MEGA):
        img, st = dev.render(cam, abi.default_render_params(W, H, spp, seed=3, kernel=kernel))
        print(name, kernel, st.paths, st.rays, float(img.mean()))
    dev.close()
host = pkg.HostScene("bvh_stress", seed=1, n=20000)
dev = pkg.DeviceScene(host.desc, device=0)
rays = np.concatenate([pkg.raysets.coherent_rays(64), pkg.raysets.incoherent_rays(4096, seed=2, extent=100.0)])
for mode in (abi.TRAVERSE_FAST, abi.TRAVERSE_EXACT):
    hits, _ = dev.trace(rays, traverse=mode)
    print("trace", mode, int((hits["prim_id"] >= 0).sum()))
host = pkg.HostScene("random_ned", seed=12345, wrap=True)
m = pkg.MultiDeviceScene(host.desc, [0, 0])
img, st = m.render(pkg.main_camera(64, 36), abi.default_render_params(64, 36, 4, seed=5))
print("multi", st.paths, st.rays)
PY
for tool in memcheck racecheck; do
  RTOW_WF_POOL=4096 timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python /tmp/san_run.py > gpurun_out/${tag}_sanitizer_${tool}.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY" gpurun_out/${tag}_sanitizer_${tool}.log
done
