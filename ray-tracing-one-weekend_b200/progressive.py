"""Progressive accumulation with checkpoint / resume.

The reference renders every frame from scratch (main.cpp:247 loops the whole image 10 000 times).  Here the per-pixel
sum and sum of squares of the sample radiance stay on the device across calls
(rtow_render_accumulate_moments_device), passes add sample indices [n, n + k) -- the Philox streams are keyed by
(pixel, sample), so K passes of k samples are the same sample set as one pass of K*k -- and the state (sums, sample
count, seed, size) can be written to disk and picked up later, e.g. for a 4096-spp 4K frame split over several jobs.

    pr = ProgressiveRender(scene, cam, 1920, 1080, seed=7)
    pr.add(64); pr.save("frame.ckpt.npz")
    pr = ProgressiveRender.load("frame.ckpt.npz", scene, cam); pr.add(64)
    mean, var_of_mean = pr.mean(), pr.variance_of_mean()

PyTorch holds the device buffers (plumbing); the kernels are the library's."""
import json

import numpy as np

from . import _abi as abi


class ProgressiveRender:
    def __init__(self, scene, cam, width, height, seed=1, max_depth=50, device=None, **params):
        import torch
        self.scene, self.cam = scene, cam
        self.width, self.height, self.seed, self.max_depth = int(width), int(height), int(seed), int(max_depth)
        self.params = dict(params)
        self.n = 0
        dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.sum = torch.zeros(self.height, self.width, 3, dtype=torch.float32, device=dev)
        self.sumsq = torch.zeros_like(self.sum)

    def add(self, samples):
        """Renders sample indices [n, n + samples) and adds them to the running moments."""
        import torch
        samples = int(samples)
        if samples <= 0:
            raise ValueError("samples must be positive")
        p = abi.default_render_params(self.width, self.height, self.n + samples, seed=self.seed, max_depth=self.max_depth,
                                      sample_begin=self.n, sample_count=samples, **self.params)
        stream = torch.cuda.current_stream(self.sum.device)
        self.scene.accumulate_moments_device(self.cam, p, self.sum.data_ptr(), self.sumsq.data_ptr(), stream.cuda_stream)
        self.n += samples
        return self.scene.stream_stats(stream.cuda_stream)

    def mean(self):
        return (self.sum / float(max(self.n, 1))).cpu().numpy()

    def variance_of_mean(self):
        """Unbiased per-pixel variance of the mean estimate: (E[x^2] - E[x]^2) / (n - 1)."""
        n = float(self.n)
        if n < 2:
            raise ValueError("needs at least two samples")
        m = self.sum.double() / n
        v = (self.sumsq.double() / n - m * m).clamp_min(0.0) / (n - 1.0)
        return v.cpu().numpy()

    def save(self, path):
        meta = dict(width=self.width, height=self.height, seed=self.seed, max_depth=self.max_depth, n=self.n, params=self.params)
        np.savez(path, sum=self.sum.cpu().numpy(), sumsq=self.sumsq.cpu().numpy(), meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8))

    @classmethod
    def load(cls, path, scene, cam, device=None):
        import torch
        z = np.load(path)
        meta = json.loads(bytes(z["meta"]).decode())
        pr = cls(scene, cam, meta["width"], meta["height"], seed=meta["seed"], max_depth=meta["max_depth"], device=device, **meta["params"])
        pr.sum.copy_(torch.from_numpy(z["sum"]))
        pr.sumsq.copy_(torch.from_numpy(z["sumsq"]))
        pr.n = int(meta["n"])
        return pr
