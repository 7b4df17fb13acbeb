import sys, numpy as np, torch
sys.path.insert(0,'.')
import __graft_entry__ as entry
pkg = entry.load_package(); abi = pkg.abi
dev_t = torch.device("cuda", 0); stream = torch.cuda.current_stream()
host = pkg.HostScene("bvh_stress", seed=12345, n=1000000)
scene = pkg.DeviceScene(host.desc, device=0)
n = 1 << 22
rays = pkg.raysets.incoherent_rays(n, seed=7, extent=100.0)
def part1by2(x):
    x = x.astype(np.uint64) & 0x3ff
    x = (x | (x << 16)) & 0x30000ff
    x = (x | (x << 8)) & 0x300f00f
    x = (x | (x << 4)) & 0x30c30c3
    x = (x | (x << 2)) & 0x9249249
    return x
def morton(rays, bits_dir):
    o = np.stack([rays['o'][:,0], rays['o'][:,1], rays['o'][:,2]], 1) if rays['o'].ndim==2 else None
    q = np.clip(((o + 100.0) / 200.0 * 1024).astype(np.int64), 0, 1023)
    m = part1by2(q[:,0]) | (part1by2(q[:,1]) << 1) | (part1by2(q[:,2]) << 2)
    d = rays['d']
    octant = ((d[:,0] < 0).astype(np.uint64)) | ((d[:,1] < 0).astype(np.uint64) << 1) | ((d[:,2] < 0).astype(np.uint64) << 2)
    if bits_dir == 'oct_hi': return (octant << 30) | m
    if bits_dir == 'oct_lo': return (m << 3) | octant
    if bits_dir == 'coarse': return ((m >> 15) << 3 | octant) << 15 | (m & 0x7fff)
    return m
print(rays.dtype)
def run(r, tag):
    d_rays = torch.from_numpy(r.view(np.uint8).reshape(len(r), -1).copy()).to(dev_t)
    d_hits = torch.empty(len(r), 48, dtype=torch.uint8, device=dev_t)
    scene.trace_device(d_rays.data_ptr(), len(r), d_hits.data_ptr(), 2, 0, stream.cuda_stream); torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(5): scene.trace_device(d_rays.data_ptr(), len(r), d_hits.data_ptr(), 2, 0, stream.cuda_stream)
    z.record(stream); torch.cuda.synchronize()
    ms = a.elapsed_time(z) / 5
    print(f"{tag}: {ms:.3f} ms, {len(r)/ms/1e3:.1f} Mrays/s", flush=True)
run(rays, "unsorted")
for mode in ("none", "oct_hi", "oct_lo", "coarse"):
    k = morton(rays, mode)
    run(np.ascontiguousarray(rays[np.argsort(k, kind='stable')]), "sorted " + mode)
