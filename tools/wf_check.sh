#!/bin/bash
# Bounds-and-invariants run of the wavefront kernels (the -DRTOW_WF_CHECK build), standing in for compute-sanitizer memcheck,
# which is closed on this GPU pool.  Build the variant on the CPU box first: tools/build_variant.sh wfcheck -DRTOW_WF_CHECK
# usage (GPU box): tools/wf_check.sh <tag>
tag=${1:-r02}
export RTOW_B200_LIB=$PWD/ray-tracing-one-weekend_b200/csrc/variants/librtow_b200_wfcheck.so
out=gpurun_out/${tag}_wf_check.txt
python - > $out 2>&1 <<'PY'
import sys, numpy as np
sys.path.insert(0, ".")
import __graft_entry__ as e
pkg = e.load_package(); abi = pkg.abi
import torch
for name, W, H, spp in (("random_ned", 200, 113, 70), ("cornell_ned", 160, 90, 20), ("cornell_smoke_ned", 160, 90, 20), ("random", 97, 61, 3)):
    host = pkg.HostScene(name, seed=12345, wrap=True)
    dev = pkg.DeviceScene(host.desc, device=0)
    cam = pkg.main_camera(W, H)
    for kw in (dict(), dict(row_begin=7, row_count=40), dict(sample_begin=1, sample_count=2), dict(traverse=abi.TRAVERSE_EXACT)):
        p = abi.default_render_params(W, H, spp, seed=3, kernel=abi.KERNEL_WAVEFRONT, **kw)
        img, st = dev.render(cam, p)
        print(name, kw, "paths", st.paths, "rays", st.rays, "mean", float(img.mean()), flush=True)
    acc = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    dev.accumulate_device_noalloc(cam, abi.default_render_params(W, H, spp, seed=3), acc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    print(name, "noalloc mean", float(acc.mean()) / spp, flush=True)
    dev.close()
PY
echo "violations: $(grep -c 'WF_CHECK violation' $out)   ok lines: $(grep -c 'WF_CHECK ok' $out)" | tee -a $out
